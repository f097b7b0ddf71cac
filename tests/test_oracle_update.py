"""Pin oracle/update.py (restatement of U SAC_CBF_CLF.update_parameters) to the 12-update trace produced by the
REFERENCE's own method (tests/golden/make_golden_update.py -> update_U.pt).  CPU, fp32: same ops in the same order, so
the bar is 1e-5 on the returned values and the multipliers, 1e-4 on the final parameters."""
import os

import torch

from oracle.update import UnicycleAgent


def _rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def test_oracle_update_matches_reference_trace(golden_dir):
    torch.set_num_threads(4)
    g = torch.load(os.path.join(golden_dir, "update_U.pt"), weights_only=False)
    ag = UnicycleAgent(g["consts"], g["B"], g["init"])
    got, ni = [], 0
    for i, u in enumerate(g["updates"]):
        nb = None
        if u % 10 == 0:
            nb, ni = g["node_batches"][ni], ni + 1
        got.append(ag.update_parameters(g["batches"][i], u, nb, g["eps"][i]))
        lam = torch.tensor([float(v) for v in ag.cs.lambda_values], dtype=torch.float64)
        assert _rel(lam, g["lambda_trace"][i]) <= 1e-5, (i, _rel(lam, g["lambda_trace"][i]))
        assert abs(ag.cs.augmented_term - g["rho_trace"][i].item()) <= 1e-12
    got = torch.tensor(got, dtype=torch.float64)
    err = (got - g["returns"]).abs().max(0).values / g["returns"].abs().max(0).values
    assert (err <= 1e-5).all(), err
    for name, sd in g["final"].items():
        for key, v in getattr(ag, name).state_dict().items():
            assert _rel(v, sd[key]) <= 1e-4, (name, key, _rel(v, sd[key]))
