"""TEST INFRASTRUCTURE — generates tests/golden/deform_*.npz by running the REAL reference
deformation module (imported from /root/reference, authoring container only) on seeded inputs.
The fixtures hold: the module's state_dict, the inputs, the 5 outputs, and the gradients of a
seeded linear loss w.r.t. every input and parameter.  Re-run: `python oracle/make_golden.py`."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)

import fdgs  # noqa: E402
from fdgs.configs import hidden_params, _kp  # noqa: E402
from fdgs.synth import perturb_deformation  # noqa: E402
import ref_deform  # noqa: E402

CASES = {
    # the DyNeRF net (arguments/dynerf/default.py:1-20: C = 16, multires [1, 2], net_width 128, all five heads) on small
    # planes: this is the shape class the tcgen05 kernels serve (the three cases below run the FP32 CUDA-core kernels or
    # the Wn = 64 instantiation).  P is not a multiple of the 128-point tile.  The loss weights of the few points whose
    # hidden pre-activations sit within 2e-5 of zero are zeroed (ref_port.relu_danger_mask) so that the fixture's
    # gradients do not hinge on a ReLU gate that FP32 re-association may flip.
    "dynerf128": dict(args=dict(kplanes_config=_kp(16, [16, 16, 16, 20]), multires=[1, 2], defor_depth=0, net_width=128,
                                no_do=False, no_dshs=False), P=2001, t=0.4375, extent=1.0, mask_danger=True),
    # debug_test.py shape (arguments/dynerf/debug_test.py:5-13), all five heads
    "tiny": dict(args=dict(kplanes_config=_kp(8, [16, 16, 16, 20]), multires=[1], defor_depth=0, net_width=32,
                           no_do=False, no_dshs=False), P=257, t=0.37, extent=1.0),
    # two levels, 3 heads, width 64, points partly OUTSIDE the aabb (border clamp -> zero coord grad)
    "small3": dict(args=dict(kplanes_config=_kp(16, [12, 10, 8, 9]), multires=[1, 2], defor_depth=1, net_width=64),
                   P=300, t=0.9967, extent=2.0),
    # M = 1 SH broadcast, t = 0
    "bcast": dict(args=dict(kplanes_config=_kp(8, [8, 8, 8, 6]), multires=[1], defor_depth=0, net_width=32,
                            no_do=False, no_dshs=False), P=64, t=0.0, extent=1.5, sh_coeffs=1),
}


def inputs(P, extent, sh_coeffs, seed):
    rng = np.random.default_rng(seed)
    f = lambda *s: torch.tensor(rng.normal(0, 1, size=s), dtype=torch.float32)
    xyz = torch.tensor(rng.uniform(-extent, extent, size=(P, 3)), dtype=torch.float32)
    return dict(xyz=xyz, scales=f(P, 3), rot=f(P, 4), opacity=f(P, 1), shs=f(P, sh_coeffs, 3))


def main(only=None):
    mod = ref_deform.load()
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    for name, case in CASES.items():
        if only and name not in only:
            continue
        args = hidden_params(**case["args"])
        torch.manual_seed(1234)
        net = mod.deform_network(args)
        perturb_deformation(net, seed=7, time_sigma=0.3)
        with torch.no_grad():   # make biases / spatial planes less regular
            for p in net.parameters():
                if p.dim() == 1:
                    p.add_(torch.randn(p.shape) * 0.05)
        net.deformation_net.set_aabb([1.2, 1.1, 1.0], [-1.0, -1.2, -1.1])
        P = case["P"]
        inp = inputs(P, case["extent"], case.get("sh_coeffs", 16), seed=11)
        leaves = {k: v.clone().requires_grad_(True) for k, v in inp.items()}
        t = torch.full((P, 1), case["t"], dtype=torch.float32)
        outs = net(leaves["xyz"], leaves["scales"], leaves["rot"], leaves["opacity"], leaves["shs"], t)
        rng = np.random.default_rng(5)
        ws = [torch.tensor(rng.normal(0, 1, size=tuple(o.shape)), dtype=torch.float32) for o in outs]
        if case.get("mask_danger"):
            import ref_port
            danger = ref_port.relu_danger_mask(net, inp["xyz"], t)
            print(name, "flip-prone points masked:", int(danger.sum()), "of", P)
            ws = [w * (~danger).to(w.dtype).reshape([-1] + [1] * (w.dim() - 1)) for w in ws]
        loss = sum((o * w).sum() for o, w in zip(outs, ws))
        loss.backward()
        blob = {}
        for k, v in net.state_dict().items():
            blob["sd/" + k] = v.detach().numpy()
        for k, v in inp.items():
            blob["in/" + k] = v.numpy()
            g = leaves[k].grad
            blob["gin/" + k] = (g if g is not None else torch.zeros_like(v)).numpy()
        for i, (o, w) in enumerate(zip(outs, ws)):
            blob[f"out/{i}"] = o.detach().numpy()
            blob[f"w/{i}"] = w.numpy()
        for k, p in net.named_parameters():
            if p.grad is not None:
                blob["gp/" + k] = p.grad.numpy()
        blob["meta/t"] = np.float32(case["t"])
        np.savez_compressed(os.path.join(out_dir, f"deform_{name}.npz"), **blob)
        print(name, "saved", sum(v.size for v in blob.values()), "values")


if __name__ == "__main__":
    main(only=sys.argv[1:] or None)
