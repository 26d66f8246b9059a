"""CPU oracle for the B200 training step — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
package, and only as the checker (or as the timed CPU baseline); the product path never routes through it.
"""
