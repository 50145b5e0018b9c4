#!/usr/bin/env python
"""Dump the init_acc_jrk forces of a seeded model (lab: compare two builds of the library, GDY_LIB=...)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from gravidy_b200.hermite4 import Hermite4B200
from gravidy_b200.plummer import plummer, default_seed
n, out = int(sys.argv[1]), sys.argv[2]
m, r, v = plummer(n, seed=default_seed(n), mass_spectrum="salpeter")
h = Hermite4B200(m, r, v)
h.init_acc_jrk()
np.save(out, h.download_forces())
