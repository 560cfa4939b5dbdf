"""openspades_b200 -- B200-native (sm_100a CUDA) implementation of OpenSpades' voxel hot path:
SW terrain pass + fog, GameMap::CastRay2, shadow bitmap + radiosity.  See DESIGN.md.

The package holds the CUDA kernels + C ABI (csrc/, built into lib/libosb200.so), the host-side C++
mirror of the reference's SWRenderer surface (host/), and this ctypes binding for tests and bench.
"""
from .binding import (LIGHT_DTYPE, MODEL_INSTANCE_DTYPE, NUM_IEEE, NUM_SSE, NUM_SSE2, RAYHIT_DTYPE, SCENE_DTYPE, Context, Group, OsbError,  # noqa: F401
                      calibrate_sse, load_library)
