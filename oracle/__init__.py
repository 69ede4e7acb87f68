"""CPU oracle for the BTS LPG decoder hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package; the product (bts_b200/) never does.
"""
