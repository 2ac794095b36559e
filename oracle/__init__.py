"""TEST INFRASTRUCTURE ONLY.

CPU restatement ("oracle") of the Perceiver IO hot path of felixyu7/perceiver-io-1.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package, and only as the checker or the timed CPU baseline.
The product path (perceiver-io-1_b200/) never imports it and has no CPU fallback.
"""
