"""nb_energy_static: the energy kernel with its scratch in a file-scope device array, for callers
that may not allocate (include/nbody_engine.h).  No counterpart in the reference; the checker is
the existing entry point nb_energy on the same bodies.  Collected last on purpose
(file name).

Device buffers of the caller are torch tensors here (the entry point itself allocates nothing);
everything goes through the C ABI (nbody.py binds nb_energy_static with ctypes)."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu

TOL = 1e-12  # the FP64 energy bound of test_gpu_parity.py at t = 0


def _device_rows(torch, pm, vel):
    dp = torch.from_numpy(np.ascontiguousarray(pm, dtype=np.float64)).cuda()
    dv = torch.from_numpy(np.ascontiguousarray(vel, dtype=np.float64)).cuda()
    return dp, dv


def _reference(nb, pm, vel):
    with nb.Engine(pm, vel, precision=nb.PREC_FP64) as e:
        return e.energy()


def test_energy_static_matches_nb_energy_twice_in_a_row(nb):
    """Two calls in a row on different sizes (the second, smaller one finds the first call's
    partials in the shared scratch array and must not read them), then the first again."""
    import torch
    g = golden("10k_bodies_g1")
    big = (g["pos_xyzm"].copy(), g["vel"].copy())
    small = (g["pos_xyzm"][:1000].copy(), g["vel"][:1000].copy())
    refs = [_reference(nb, *big), _reference(nb, *small)]
    out = torch.full((3, 2), float("nan"), dtype=torch.float64, device="cuda")
    bufs = [_device_rows(torch, *big), _device_rows(torch, *small)]
    torch.cuda.synchronize()
    for k, which in enumerate([0, 1, 0]):
        dp, dv = bufs[which]
        nb.energy_static(dp.data_ptr(), dv.data_ptr(), dp.shape[0], out[k].data_ptr())
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    for k, which in enumerate([0, 1, 0]):
        ke, pe = refs[which]
        assert abs(got[k, 0] - ke) <= TOL * abs(ke), (k, got[k, 0], ke)
        assert abs(got[k, 1] - pe) <= TOL * abs(pe), (k, got[k, 1], pe)
    # same kernel, same summation order: in fact the same bits
    assert got[0, 0] == refs[0][0] and got[0, 1] == refs[0][1]
    assert np.array_equal(got[0], got[2])


def test_energy_static_honours_G_stream_and_sizes(nb):
    import torch
    g = golden("1k_bodies")
    for n in (1, 2, 255, 513):
        pm, vel = g["pos_xyzm"][:n].copy(), g["vel"][:n].copy()
        with nb.Engine(pm, vel, precision=nb.PREC_FP64, G=2.5) as e:
            ke, pe = e.energy()
        dp, dv = _device_rows(torch, pm, vel)
        out = torch.zeros(2, dtype=torch.float64, device="cuda")
        s = torch.cuda.Stream()
        torch.cuda.synchronize()
        nb.energy_static(dp.data_ptr(), dv.data_ptr(), n, out.data_ptr(), G=2.5, stream=s.cuda_stream)
        s.synchronize()
        got = out.cpu().numpy()
        assert abs(got[0] - ke) <= TOL * max(abs(ke), 1e-300)
        assert abs(got[1] - pe) <= TOL * max(abs(pe), 1e-300)


def test_energy_static_refuses_what_does_not_fit(nb):
    import torch
    out = torch.zeros(2, dtype=torch.float64, device="cuda")
    dummy = torch.zeros(8, dtype=torch.float64, device="cuda")
    with pytest.raises(nb.NbodyError) as ei:
        nb.energy_static(dummy.data_ptr(), dummy.data_ptr(), nb.ENERGY_STATIC_MAX_BODIES + 1, out.data_ptr())
    assert ei.value.code == -1  # NB_E_INVALID
    assert "NB_ENERGY_STATIC_MAX_BODIES" in str(ei.value)
    with pytest.raises(nb.NbodyError):
        nb.energy_static(None, None, 4, out.data_ptr())
    # the refusals left the scratch path usable
    g = golden("1k_bodies")
    dp, dv = _device_rows(torch, g["pos_xyzm"], g["vel"])
    nb.energy_static(dp.data_ptr(), dv.data_ptr(), dp.shape[0], out.data_ptr())
    torch.cuda.synchronize()
    ke, pe = _reference(nb, g["pos_xyzm"], g["vel"])
    got = out.cpu().numpy()
    assert got[0] == ke and got[1] == pe
