"""Pin oracle/constraints.py to the golden vectors produced by the REFERENCE's own get_policy_loss_2 and
DynamicsModel (tests/golden/make_golden_constraints.py, run in the build container)."""
import os

import pytest
import torch

from oracle import constraints as OC
from oracle import nlbac_ref
from constraint_cases import build, run_backup_sequence, run_sequence


def _close(a, b, rtol=2e-5, atol=1e-6):
    scale = max(b.abs().max().item(), 1e-12)
    return (a - b).abs().max().item() <= rtol * scale + atol


@pytest.mark.parametrize("tag", ["U", "C", "BU", "P", "BP"])
def test_policy_loss_2_matches_reference(tag, golden_dir):
    g = torch.load(os.path.join(golden_dir, "constraints_%s.pt" % tag), weights_only=False)
    nets = build(g, nlbac_ref, "cpu")
    losses, grads, st = run_sequence(g, nets, OC, "cpu")
    for got, want in zip(losses, g["losses"]):
        assert _close(got, want.reshape(-1)), (got, want)
    for got, want in zip(grads, g["grad_action"]):
        assert _close(got, want, rtol=1e-4), (got - want).abs().max()
    assert torch.allclose(torch.tensor([float(v) for v in st.lambda_values]), torch.tensor(g["lambda_after"]), rtol=1e-5)
    assert abs(st.augmented_term - g["augmented_term_after"]) < 1e-12


@pytest.mark.parametrize("tag", ["U", "C", "P"])
def test_backup_policy_loss_2_matches_reference(tag, golden_dir):
    g = torch.load(os.path.join(golden_dir, "constraints_backup_%s.pt" % tag), weights_only=False)
    losses, grads, st = run_backup_sequence(g, build(g, nlbac_ref, "cpu"), OC, "cpu")
    for got, want in zip(losses, g["losses"]):
        assert _close(got, want.reshape(-1)), (got, want)
    for got, want in zip(grads, g["grad_action"]):
        assert _close(got, want, rtol=1e-4, atol=1e-10), (got - want).abs().max()
    assert torch.allclose(torch.tensor([float(v) for v in st.backup_lambda_values]), torch.tensor(g["lambda_after"]),
                          rtol=1e-5)
    rho = st.backup_augmented_term if tag == "P" else st.augmented_term
    assert abs(rho - g["augmented_term_after"]) < 1e-12


@pytest.mark.parametrize("tag,env", [("U", "Unicycle"), ("C", "SimulatedCars"), ("BU", "Unicycle")])
def test_get_state_matches_reference(tag, env, golden_dir):
    g = torch.load(os.path.join(golden_dir, "constraints_%s.pt" % tag), weights_only=False)
    assert torch.equal(OC.get_state(env, g["obs"]), g["state_from_obs"])


def test_get_obs_and_grad_match_reference(golden_dir):
    g = torch.load(os.path.join(golden_dir, "constraints_BU.pt"), weights_only=False)
    s = g["state"].clone().requires_grad_(True)
    obs = OC.get_obs("Unicycle", s)
    assert _close(obs.detach(), g["obs"])
    (obs * g["obs_cot"]).sum().backward()
    assert _close(s.grad, g["grad_state_get_obs"], rtol=1e-4)
    p = torch.load(os.path.join(golden_dir, "constraints_Pmaps.pt"), weights_only=False)
    s = p["state7"].clone().requires_grad_(True)
    obs = OC.get_obs("Pvtol", s)
    assert _close(obs.detach(), p["obs"])
    (obs * p["obs_cot"]).sum().backward()
    assert _close(s.grad, p["grad_state_get_obs"], rtol=1e-4)
    st, dyn = OC.get_state("Pvtol", p["obs"])
    assert torch.equal(st, p["state_from_obs"]) and torch.equal(dyn, p["state_dyn_from_obs"])
    c = torch.load(os.path.join(golden_dir, "constraints_C.pt"), weights_only=False)
    assert _close(OC.get_obs("SimulatedCars", c["state"]), c["obs"])
