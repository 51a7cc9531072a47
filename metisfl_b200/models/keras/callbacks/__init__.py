"""Import-path parity (reference ``metisfl/models/keras/callbacks/``)."""
