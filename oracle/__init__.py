"""CPU checkers for aircraft_b200 -- TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke()
and bench.py's CPU-baseline legs may import this package; the product never does."""
