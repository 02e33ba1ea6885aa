"""Config-5 top-2 search once (development aid for ncu)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from orb_slam3_detailed_comments_kor_b200 import DescriptorDatabase, synth
db, q = synth.descriptors(5, 1 << 20, 4096)
m = DescriptorDatabase(1 << 20, 4096); m.set(db)
dq = torch.from_numpy(q).cuda(); out = torch.zeros((4096, 4), dtype=torch.int32, device="cuda")
s = torch.cuda.Stream()
for _ in range(3):
    m.top2_device(dq.data_ptr(), 4096, 50, 0.9, False, out.data_ptr(), s.cuda_stream)
torch.cuda.synchronize(); print("accepted", int(out[:, 3].sum()))
