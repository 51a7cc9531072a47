"""Import-path parity with reference ``metisfl/models/keras/optimizers/fed_prox.py``."""
from metisfl_b200.optim.fed_prox import FedProx  # noqa: F401
