"""Operators: ctypes binding of the native sm_100a library (``native``), layout packers (``layout``) and thin
tensor-level wrappers (``flat``, ``conv``)."""
