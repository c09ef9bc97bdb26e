"""Hand-written sm_100a ops (ctypes front-ends) and their CPU oracle."""
