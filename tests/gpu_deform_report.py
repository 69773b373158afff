"""Diagnostic: fused deformation kernels vs golden fixtures (real reference) and vs the torch oracle."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import fdgs  # noqa: E402
from fdgs.configs import hidden_params, _kp  # noqa: E402
from fdgs.deformation import deform_network  # noqa: E402
import util  # noqa: E402

CASES = {
    "tiny": dict(kplanes_config=_kp(8, [16, 16, 16, 20]), multires=[1], defor_depth=0, net_width=32, no_do=False, no_dshs=False),
    "small3": dict(kplanes_config=_kp(16, [12, 10, 8, 9]), multires=[1, 2], defor_depth=1, net_width=64),
    "bcast": dict(kplanes_config=_kp(8, [8, 8, 8, 6]), multires=[1], defor_depth=0, net_width=32, no_do=False, no_dshs=False),
}


def run_case(name):
    z = np.load(os.path.join(ROOT, "tests", "golden", f"deform_{name}.npz"))
    args = hidden_params(**CASES[name])
    net = deform_network(args)
    sd = {k[3:]: torch.tensor(z[k]) for k in z.files if k.startswith("sd/")}
    net.deformation_net.grid.set_aabb(sd["deformation_net.grid.aabb"][0].tolist(), sd["deformation_net.grid.aabb"][1].tolist())
    net.load_state_dict(sd)
    net = net.cuda()
    inp = {k[3:]: torch.tensor(z[k]).cuda().requires_grad_(True) for k in z.files if k.startswith("in/")}
    t = float(z["meta/t"])
    P = inp["xyz"].shape[0]
    tt = torch.full((P, 1), t, device="cuda")
    outs = net(inp["xyz"], inp["scales"], inp["rot"], inp["opacity"], inp["shs"], tt)
    loss = sum((o * torch.tensor(z[f"w/{i}"]).cuda()).sum() for i, o in enumerate(outs))
    loss.backward()
    torch.cuda.synchronize()
    print(f"== deform {name}: P={P}")
    for i, o in enumerate(outs):
        print(f"   out[{i}] rel err {util.rel_err(o, z[f'out/{i}']):.3e}")
    for k in z.files:
        if k.startswith("gin/"):
            g = inp[k[4:]].grad
            print(f"   grad in {k[4:]}: {util.rel_err(g if g is not None else torch.zeros(1), z[k]):.3e}")
    worst = 0
    for k, p in net.named_parameters():
        key = "gp/" + k
        if key in z.files:
            e = util.rel_err(p.grad, z[key]) if p.grad is not None else float("nan")
            worst = max(worst, e) if e == e else float("nan")
            if e != e or e > 1e-4:
                print(f"   grad param {k}: {e:.3e}")
    print(f"   worst param-grad rel err {worst:.3e}")


if __name__ == "__main__":
    for n in CASES:
        run_case(n)
