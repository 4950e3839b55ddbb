"""oracle/ — CPU checker for the ORB + Hamming-kNN front-end (TEST INFRASTRUCTURE ONLY).

Importable only from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs.  The product package (keyframe-based-slam_b200/) never imports it.
"""
