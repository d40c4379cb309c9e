"""Long-horizon parity: the fused step against the fp32 restatement (oracle/restated.py evaluated on the GPU, TF32
off) over thousands of steps from the same initial state and batches.  Two implementations of the same fp32
algorithm drift apart chaotically, so this is evidence for profiles/, not a gate: it reports how the per-step loss /
FVU / L0 deviations develop.
    python scripts/long_run_check.py [steps] > gpurun_out/long_run.json"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import sparse_coding__b200 as sc  # noqa: E402
from conftest import make_sparse_data  # noqa: E402
from oracle import restated as R  # noqa: E402  (test infrastructure; this script is a check, not product code)


def main():
    steps = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
    torch.backends.cuda.matmul.allow_tf32 = False
    M, B, d, N = 4, 1024, 256, 1024
    out = {"shape": {"M": M, "B": B, "d": d, "N": N}, "steps": steps}
    for fmt in ("auto", "tf32"):
        torch.manual_seed(0)
        models = [sc.FunctionalSAE.init(d, N, float(l1), device="cuda") for l1 in np.logspace(-3.5, -2.5, M)]
        ens = sc.FunctionalEnsemble(models, sc.FunctionalSAE, sc.torchopt_compat.adam, {"lr": 1e-3}, device="cuda")
        ens.operand_format = fmt
        orc = R.EnsembleOracle({k: v.clone() for k, v in ens.params.items()}, {k: v.clone() for k, v in ens.buffers.items()},
                               tied=False, lr=1e-3)
        data = make_sparse_data(64 * B, d, 2 * d, 8, seed=1).cuda()
        g = torch.Generator().manual_seed(2)
        rows = []
        for s in range(steps):
            x = data[torch.randint(0, data.shape[0], (B,), generator=g).cuda()]
            lg, ag = ens.step_batch(x)
            lr_, ar = orc.step_batch(x)
            if s % 100 == 99 or s == steps - 1:
                dl = float(((lg["loss"] - lr_["loss"]).abs() / lr_["loss"].abs()).max())
                df = float(((lg["l_reconstruction"] - lr_["l_reconstruction"]).abs() / lr_["l_reconstruction"].abs()).max())
                l0g = (ag["c"] != 0).float().sum(-1).mean(-1)
                l0r = (ar["c"] != 0).float().sum(-1).mean(-1)
                d0 = float(((l0g - l0r).abs() / l0r.clamp_min(1.0)).max())
                rows.append({"step": s + 1, "loss_rel": dl, "fvu_rel": df, "l0_rel": d0,
                             "loss_ref": [round(float(v), 6) for v in lr_["loss"]], "l0_ref": [round(float(v), 1) for v in l0r]})
        out[ens.last_operand_format] = {"checkpoints": rows,
                                        "max_loss_rel": max(r["loss_rel"] for r in rows), "max_fvu_rel": max(r["fvu_rel"] for r in rows),
                                        "max_l0_rel": max(r["l0_rel"] for r in rows)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
