"""oracle/ — TEST INFRASTRUCTURE ONLY. CPU restatement of the reference's hot-path algorithms
(pebble_oracle.c) and a ctypes binding to the real reference built from /root/reference (oracle/_ref).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package; the product (pebblesdr_b200) never does."""
