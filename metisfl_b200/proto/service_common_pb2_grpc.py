"""No services are declared in service_common.proto; kept for import parity with protoc output."""
