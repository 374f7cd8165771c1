"""Allocation-free entry point epp_check_poly_trajectories_static_dev (include/epp_c_api.h; no counterpart in the
reference) against the existing epp_check_poly_trajectories_dev, both
called through the C ABI.  Integer / bit outputs: the comparison is bit-exact, like every other GPU test here.

This file is named so that pytest collects it LAST: a failure here must not hide
the parity tests."""
import ctypes as C

import numpy as np
import pytest

import epp_b200
from epp_b200 import worlds

pytestmark = pytest.mark.gpu

MAX_SAMPLES = 1024
MIN_DISTANCE = 0.1


def _world(config):
    g, o, lo, hi = worlds.world_256()
    w = epp_b200.World(config)
    w.parse_gates_and_obstacles(g, o)
    w.commit()  # the allocation-free call never commits: the world is resident from here on
    return w, lo, hi


def _batch(n_traj, lo, hi, seed, dev):
    import torch
    coeffs, T, dt = worlds.rest_to_rest_trajectories(n_traj, 8, lo, hi, seed=seed, n_samples=1000, max_hop=3.0)
    return tuple(torch.from_numpy(np.ascontiguousarray(x)).to(dev) for x in (coeffs, T, dt))


def _outputs(n_traj, dev, fill):
    """Caller-owned result buffers, pre-filled with garbage the call must overwrite."""
    import torch
    valid = torch.full((epp_b200.mask_words(n_traj),), fill, dtype=torch.int32, device=dev)
    smask = torch.full((n_traj, MAX_SAMPLES // 32), fill, dtype=torch.int32, device=dev)
    ns = torch.full((n_traj,), fill, dtype=torch.int32, device=dev)
    first = torch.full((n_traj,), fill, dtype=torch.int64, device=dev)
    return valid, smask, ns, first


def _same(got, want, n_traj):
    import torch
    torch.cuda.synchronize()
    names = ("trajectory validity words", "sample mask", "sample counts", "first invalid row")
    for g, w, name in zip(got, want, names):
        if g is None:
            continue
        assert g.shape == w.shape and g.dtype == w.dtype, name
        assert torch.equal(g, w), "%s differ between the static-scratch and the caller-scratch entry point" % name
    bits = np.unpackbits(want[0].cpu().numpy().view(np.uint8), bitorder="little")[:n_traj]
    return int(bits.sum())


def test_static_scratch_matches_caller_scratch_two_calls_in_a_row(config):
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    w, lo, hi = _world(config)

    # call 1: 300 trajectories; call 2: a DIFFERENT, smaller batch, so the scratch words of call 1 (some of them
    # lowered to a row number by its invalid trajectories) are what call 2 finds in the array
    batches = [(300, worlds.SEED_TRAJ), (170, worlds.SEED_TRAJ + 17)]
    n_valid = []
    for k, (n_traj, seed) in enumerate(batches):
        coeffs, T, dt = _batch(n_traj, lo, hi, seed, dev)
        want = w.check_poly_trajectories_dev(coeffs, T, dt, MIN_DISTANCE, max_samples=MAX_SAMPLES)
        got = _outputs(n_traj, dev, 0x5A5A5A5A if k == 0 else -1)
        torch.cuda.synchronize()
        held = (torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev))
        w.check_poly_trajectories_static_dev(coeffs, T, dt, MIN_DISTANCE, MAX_SAMPLES, *got)
        assert (torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)) == held
        n_valid.append(_same(got, want, n_traj))
        assert 0 < n_valid[-1] < n_traj  # both verdicts occur: stale "invalid" words would show

    # the same batch again, twice in a row, without the optional outputs (d_first_invalid NULL: the case the
    # caller-scratch entry point cannot serve), results unchanged
    n_traj, seed = batches[0]
    coeffs, T, dt = _batch(n_traj, lo, hi, seed, dev)
    want = w.check_poly_trajectories_dev(coeffs, T, dt, MIN_DISTANCE, max_samples=MAX_SAMPLES)
    for _ in range(2):
        valid = torch.full((epp_b200.mask_words(n_traj),), 0x33333333, dtype=torch.int32, device=dev)
        w.check_poly_trajectories_static_dev(coeffs, T, dt, MIN_DISTANCE, MAX_SAMPLES, valid)
        assert _same((valid, None, None, None), want, n_traj) == n_valid[0]


def test_static_scratch_rejects_batches_beyond_its_maximum(config):
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    w, lo, hi = _world(config)
    coeffs, T, dt = _batch(4, lo, hi, worlds.SEED_TRAJ, dev)
    valid = torch.zeros(8, dtype=torch.int32, device=dev)
    L = epp_b200.lib()
    launches = epp_b200.kernel_launch_count()
    too_many = epp_b200.STATIC_MAX_TRAJECTORIES + 1
    rc = L.epp_check_poly_trajectories_static_dev(w.handle, coeffs.data_ptr(), T.data_ptr(), dt.data_ptr(), too_many, 8, 10,
                                                  MIN_DISTANCE, MAX_SAMPLES, valid.data_ptr(), None, None, None,
                                                  C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 6  # EPP_ERR_INVALID_ARG
    message = L.epp_last_error().decode()
    assert "EPP_STATIC_MAX_TRAJECTORIES" in message and str(epp_b200.STATIC_MAX_TRAJECTORIES) in message
    assert epp_b200.kernel_launch_count() == launches  # nothing was launched
    torch.cuda.synchronize()
    assert int(valid.abs().sum().item()) == 0

    # an uncommitted world is an error, not a hidden commit (which would allocate)
    fresh = epp_b200.World(config)
    with pytest.raises(epp_b200.EppError) as err:
        fresh.check_poly_trajectories_static_dev(coeffs, T, dt, MIN_DISTANCE, MAX_SAMPLES, valid)
    assert err.value.status == 7  # EPP_ERR_NOT_COMMITTED
