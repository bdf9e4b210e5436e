import os, sys, subprocess
CHILD = r'''
import sys, torch, numpy as np
sys.path.insert(0, "/root/repo")
from aircraft_b200 import Aircraft
from aircraft_b200.config import TRIM_COM
from aircraft_b200.workloads import make_states, make_control_sequences
ac = Aircraft.from_problem(com=TRIM_COM, device=0, physical_integration_substeps=1, normalise=True)
B, H = 512, 20
x0, U = make_states(B, seed=45), make_control_sequences(B, H, seed=45, dtype=np.float32)
X0 = torch.tensor(x0.T.copy(), dtype=torch.float32, device="cuda"); UU = torch.tensor(U.transpose(1, 2, 0).copy(), device="cuda")
a = ac.rollout(X0, UU, 0.01, layout="soa", normalise=True, cost=True)
torch.save((a.final.cpu(), a.cost.cpu()), sys.argv[1])
one = ac.state_update(X0, UU[0], 0.01)
torch.save(one.cpu(), sys.argv[1] + ".step")
'''
outs = {}
for coop in ("0", "2", "4", "8"):
    f = "/tmp/coop_%s.pt" % coop
    r = subprocess.run([sys.executable, "-c", CHILD, f], env=dict(os.environ, AIRC_ROLLOUT_COOP=coop), capture_output=True, text=True)
    if r.returncode: print(r.stderr[-500:])
import torch
ref = torch.load("/tmp/coop_8.pt")
for coop in ("0", "2", "4"):
    o = torch.load("/tmp/coop_%s.pt" % coop)
    d = (o[0] - ref[0]).abs()
    print("L=%s vs L=8: final equal %s  max abs diff %.3e  n differing %d  cost equal %s" % (coop, torch.equal(o[0], ref[0]), float(d.max()), int((d > 0).sum()), torch.equal(o[1], ref[1])))
