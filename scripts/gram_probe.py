#!/usr/bin/env python
"""Registers a 1024-column x 262144-row Gaussian data set once (one launch of the
Gram kernel) -- a short command for ncu."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bnlearn_b200 import Engine
rng = np.random.default_rng(0)
x = rng.normal(size=(1024, 262144))
with Engine() as e:
    d = e.register(list(x))
    print(e.last_call_profile())
