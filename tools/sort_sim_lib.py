# helpers shared by the warp-model experiments (see tools/sort_sim.py)
