"""TEST INFRASTRUCTURE ONLY -- stand-in: droid_slam/geom/graph_utils.py:7 imports two dataset helpers that the hot path
never calls (training only); the reference's data_readers package pulls in dataset code that is out of scope here."""
