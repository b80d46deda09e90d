"""TEST INFRASTRUCTURE ONLY -- empty stand-in: droid_slam/factor_graph.py:5 imports matplotlib.pyplot and never uses it."""
