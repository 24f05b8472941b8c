"""Latency block of bench.py alone (the reference's call pattern): python tools/exp_latency2.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
import orb_slam2_b200 as gx
print(json.dumps(bench.latency_block(gx, torch, 0), indent=1))
