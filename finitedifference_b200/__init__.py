"""B200-native (sm_100a) implicit 2D inviscid-Burgers stepper: FOM, LSPG PROM, ECSW HPROM.

Drop-in for the hot path of SADPR/FiniteDifference's ``hypernet2D.py``; see
``finitedifference_b200.hypernet2D``.  Importing the package does not need a GPU; running a
solver does (there is no CPU path).
"""
__version__ = "0.1.0"
