"""CPU oracle for the lattice filter / mean-field CRF hot path.

TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this package; the product
package (depth-estimation_b200/) never does.
"""
