"""TEST INFRASTRUCTURE (not product): the CPU oracle of the fast-binless path. Only tests/, __graft_entry__.smoke() and
bench.py's CPU legs import this package; nothing under binless_b200/ does."""
