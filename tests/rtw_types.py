import numpy as np

RAY_DTYPE = np.dtype([("origin", "<f4", 3), ("tmin", "<f4"), ("dir", "<f4", 3), ("tmax", "<f4"), ("time", "<f4"),
                      ("gather_time", "<f4"), ("seed", "<u4"), ("pad", "<u4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("prim", "<i4")])

# (scene id, Nx, Ny) used by the parity tests: small enough for the CPU oracle
SCENE_CASES = [(0, 96, 96), (1, 96, 54), (2, 96, 54), (3, 64, 64), (4, 64, 64)]
