"""No services are declared in metis.proto; kept for import parity with protoc output."""
