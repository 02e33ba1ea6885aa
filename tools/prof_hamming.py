#!/usr/bin/env python3
"""Smallest program that runs the 4096 x 2^20 top-2 search a few times (for ncu captures under gpurun)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from orb_slam3_detailed_comments_kor_b200 import DescriptorDatabase, synth

nq, ndb = 4096, 1 << 20
db, q = synth.descriptors(5, ndb, nq)
shard = DescriptorDatabase(ndb, nq)
shard.set(db)
d_q = torch.from_numpy(q).cuda()
d_out = torch.zeros((nq, 4), dtype=torch.int32, device="cuda")
st = torch.cuda.Stream()
torch.cuda.synchronize()
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    shard.top2_device(d_q.data_ptr(), nq, 50, 0.9, False, d_out.data_ptr(), st.cuda_stream)
st.synchronize()
print("ok: accepted", int(d_out[:, 3].sum().item()))
