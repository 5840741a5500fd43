"""CPU oracle for the fo-rma cpu_ray_tracer hot path — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package; the product (fo-rma_b200/) never does.  See forma_oracle.cpp for the pin status.
"""
