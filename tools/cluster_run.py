"""World::ClusterizeBlocks of every solid voxel of Title.vxl above z = 48 (the bench's extra.links.clusterize case), three
calls; the command profiled by tools/gpu_cluster_ncu.sh."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import openspades_b200 as osb
from openspades_b200 import workloads
from openspades_b200.hostlib import VoxelMap

m = VoxelMap.load_file(workloads.TITLE_VXL)
solid = m.solid
xs, ys = np.nonzero(solid & np.uint64((1 << 48) - 1))
cells = []
for z in range(48):
    sel = ((solid[xs, ys] >> np.uint64(z)) & np.uint64(1)).astype(bool)
    cells.append(np.stack([xs[sel], ys[sel], np.full(int(sel.sum()), z)], 1))
cells = np.ascontiguousarray(np.concatenate(cells).astype(np.int16))
with osb.Context(device=0, numerics=osb.NUM_IEEE) as ctx:
    for _ in range(3):
        k, lab = ctx.clusterize_blocks(cells)
print(len(cells), "cells,", k, "pieces")
