"""TEST INFRASTRUCTURE ONLY.

CPU/GPU restatement of the reference's attention math and a launcher for the reference's own
kernel.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this package; the product (flash_attention_cute_b200/) never does.
"""
