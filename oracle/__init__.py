"""CPU oracle for the relay-audio ingest path — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package; the product (loqa_hub_b200) never does.
"""
