"""No services are declared in model.proto; kept for import parity with protoc output."""
