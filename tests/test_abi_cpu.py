"""CPU checks of the boundary: the C-ABI library loads, exports every symbol include/nlbac.h declares,
and the host-side mirror keeps the reference's call surface.  No compute calls (no GPU here)."""
import ctypes

import pytest
import torch


@pytest.fixture(scope="module")
def pkg():
    import nlbac_b200
    nlbac_b200.build()
    return nlbac_b200


def test_library_exports_every_declared_symbol(pkg):
    L = pkg._lib.lib()
    names = pkg._lib.declared_symbols()
    assert len(names) >= 10
    for n in names:
        assert hasattr(L, n), n
    assert L.nlbac_abi_version() == 2


def test_descriptor_queries_match_reference_shapes(pkg):
    L = pkg._lib.lib()
    for ctor, D, n_param in (((3, 3, 6), 5, 52209), ((6, 6, 12), 8, 53718), ((12, 10), 12, 9802)):
        m = pkg.NeuralODEModel(*ctor)
        assert sum(p.numel() for p in m.parameters()) == n_param
        desc, params = pkg.describe(m, D)
        assert L.nlbac_node_supported(ctypes.byref(desc), 0) == 0
        assert L.nlbac_node_param_count(ctypes.byref(desc)) == n_param
        assert sum(p.numel() for p in params) == n_param
    bad = pkg._lib.NodeDesc(0, 4, 2, 0, 100, 5, 4)
    assert L.nlbac_node_supported(ctypes.byref(bad), 0) != 0
    assert b"no compiled kernel instance" in L.nlbac_last_error()


def test_state_dict_keys_are_the_reference_keys(pkg):
    m = pkg.NeuralODEModel(3, 3, 6)
    assert list(m.state_dict()) == ["f_net.%d.%s" % (i, k) for i in (0, 2, 4, 6, 8) for k in ("weight", "bias")] + \
        ["g_net.%d.%s" % (i, k) for i in (0, 2, 4, 6) for k in ("weight", "bias")]
    c = pkg.NeuralODEModel(12, 10)
    assert list(c.state_dict()) == ["net.%d.%s" % (i, k) for i in (0, 2, 4, 6) for k in ("weight", "bias")]


def test_odeint_argument_errors_without_gpu(pkg):
    m = pkg.NeuralODEModel(3, 3, 6)
    y0 = torch.zeros(2, 5)
    with pytest.raises(ValueError):
        pkg.odeint(m, y0, torch.tensor([0.0, 1.0]), method="bogus")
    with pytest.raises(RuntimeError):          # CPU tensors are refused: no fallback
        pkg.odeint(m, y0, torch.tensor([0.0, 1.0]), method="euler")
    with pytest.raises(NotImplementedError):
        pkg.odeint(m, (y0, y0), torch.tensor([0.0, 1.0]), method="euler")


def test_forward_api_matches_oracle_rhs(pkg):
    """NeuralODEModel.forward(t, s) (API compatibility) equals the oracle restatement of the reference forward."""
    from oracle import nlbac_ref
    torch.manual_seed(3)
    for ctor, ref, D in (((3, 3, 6), nlbac_ref.AffineNODE(3, 3, 6), 5), ((12, 10), nlbac_ref.BlackBoxNODE(12, 10), 12)):
        m = pkg.NeuralODEModel(*ctor)
        m.load_state_dict(ref.state_dict())
        s = torch.randn(7, D)
        assert torch.allclose(m(None, s), ref(None, s), atol=1e-6)


def test_import_shim_aliases_submodules():
    """``from nlbac_b200.odeint import ...`` must reuse the already-loaded submodule: a second copy would carry its own
    caches and rebind ``nlbac_b200.odeint`` (the function) to the module (regression: round 2)."""
    import sys
    import nlbac_b200
    from nlbac_b200.odeint import describe            # noqa: F401
    from nlbac_b200 import tail, parallel, dopri5      # noqa: F401
    assert callable(nlbac_b200.odeint) and not isinstance(nlbac_b200.odeint, type(sys))
    long_name = nlbac_b200.__name__
    for sub in ("odeint", "tail", "constraints", "agent", "parallel", "dopri5", "_lib"):
        assert sys.modules["nlbac_b200." + sub] is sys.modules[long_name + "." + sub], sub
