#!/usr/bin/env python
"""Two rayca_cuda_scene_create calls on the atrium and nothing else: the launch list of a scene build
(ncu --metrics gpu__time_duration.sum ... python tools/build_only.py)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rayca_b200 import procedural as P, renderer as R

ctx = R.Context(0)
desc = P.atrium(262_267).to_desc()
for _ in range(2):
    R.DeviceScene(ctx, desc).close()
