#!/bin/bash
# force_cmp.sh libA libB n...: forces of two library builds side by side
A=$1; B=$2; shift 2
for n in "$@"; do
  GDY_LIB=$PWD/$A python tools/lab/force_dump.py $n /tmp/fa.npy
  GDY_LIB=$PWD/$B python tools/lab/force_dump.py $n /tmp/fb.npy
  python - <<PY
import numpy as np
a, b = np.load("/tmp/fa.npy"), np.load("/tmp/fb.npy")
bad = np.where((a != b).any(axis=1))[0]
print("n=$n differing particles:", len(bad), "of", len(a), "first", bad[:20], "last", bad[-5:])
if len(bad):
    k = bad[0]
    print(" a", a[k]); print(" b", b[k]); print(" ratio", b[k] / a[k])
    rel = np.linalg.norm(a[:, :3] - b[:, :3], axis=1) / np.linalg.norm(a[:, :3], axis=1)
    print(" max rel", rel.max(), "median rel of bad", np.median(rel[bad]))
PY
done
