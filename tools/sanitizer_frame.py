"""One small LBVH frame for compute-sanitizer: scene 4 (3410 primitives, persistent walkers, NEE), two waves
forced, the drain tail included; then the by-value small scene with the fused NEE; then the 2-context path."""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rtw = importlib.import_module("raytracing-weekend_b200")
d = rtw.Director(0)
d.init(96, 64, 4)
d.set_option("waves", 2)
d.set_option("pool_paths", 4096)
d.createScene(4)
d.renderFrame(4)
print("scene 4:", d.stats()["paths"], "paths,", d.stats()["kernel_launches"], "launches")
d.createScene(0)
d.renderFrame(4)
print("scene 0:", d.stats()["paths"], "paths")
d.createScene(1)
d.renderFrame(2)
print("scene 1:", d.stats()["paths"], "paths")
d.destroy()
m = rtw.MultiDirector([0, 0])
m.init(64, 48, 3)
m.createScene(3)
m.renderFrame(3)
print("multi:", m.stats()["paths"], "paths")
m.destroy()
