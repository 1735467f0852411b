"""timing only: fused2 vs layer-wise at M = 1M rows (K1b development)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import tools.chain_bwd_check as C
C.run(1048576, True, True, time_it=True)
