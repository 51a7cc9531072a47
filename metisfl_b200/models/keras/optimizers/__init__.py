"""Import-path parity (reference ``metisfl/models/keras/optimizers/``); the implementation is framework-neutral."""
