"""oracle/ -- TEST INFRASTRUCTURE ONLY: CPU restatement of the reference path (port.c), the recipe that compiles the
unmodified reference (oracle/_ref), and the parity checker.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this package; meshlib_b200/ never does."""
