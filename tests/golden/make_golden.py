"""Copies the reference's golden (sqrt(NUM_DOFS), energy error) rows that are small enough for the CPU suite
into tests/golden/tutorial_01_uniform_mesh_error.json.  Run in the build container (reads /root/reference)."""
import json
import os

SRC = "/root/reference/Tutorials/tutorial_01_poisson_sharp_circular_wave/convergence_plot/uniform_mesh_error_order_%d"
out = {}
for order in (1, 2, 4):
    rows = []
    for line in open(SRC % order):
        a, b = line.split()
        if float(a) <= 100:
            rows.append([float(a), float(b)])
    out[str(order)] = rows
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "tutorial_01_uniform_mesh_error.json"), "w"), indent=1)
