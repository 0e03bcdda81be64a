"""CPU oracle for the hot path -- TEST INFRASTRUCTURE, not product code.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs may
import from here.  neuralangelo_b200/ never does.  See oracle/README.md.
"""
