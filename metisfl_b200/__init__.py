"""metisfl_b200 — a Blackwell-native federated-learning framework with MetisFL's capabilities.

Layout: ``proto`` wire contract, ``controller``/``learner``/``driver`` services,
``models`` + ``optim`` learner framework layer, ``ops`` sm_100a kernels, ``parallel`` peer-memory
fabric and the on-box federation engine, ``encryption`` CKKS secure aggregation, ``utils``.
"""
__version__ = "0.1.0"
