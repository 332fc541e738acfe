"""b200cs_mappm_static: the no-allocation variant of b200cs_mappm (include/b200cs.h).  No counterpart
in the reference; added for callers that cannot cudaMalloc.
Collected last on purpose.  Calls go through the C-ABI with ctypes;
the outputs must equal those of b200cs_mappm (same kernels, so bit for bit -- the tolerance of
test_mappm_all_kord_bitexact) on two calls in a row that share the static workspace."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu

STATIC_MAX_BYTES = 16 * 1024 * 1024  # B200CS_MAPPM_STATIC_MAX_BYTES


def _columns(rng, ncol, km, kn):
    ps1 = 100000.0 + 2000.0 * rng.standard_normal(ncol)
    ps2 = ps1 + 300.0 * rng.standard_normal(ncol)
    pe1 = 300.0 + (ps1[:, None] - 300.0) * (np.linspace(0, 1, km + 1) ** 1.5)[None]
    pe2 = 300.0 + (ps2[:, None] - 300.0) * (np.linspace(0, 1, kn + 1) ** 1.3)[None]
    return pe1.astype(np.float32), pe2.astype(np.float32)


def _call(lib, name, pe1, q, pe2, iv, kord, ws=None):
    """one C-ABI call on [ncol, nz] device arrays; returns (rc, out)"""
    ncol, km = q.shape
    kn = pe2.shape[1] - 1
    out = torch.full((ncol, kn), float("nan"), dtype=torch.float32, device="cuda")
    args = [pe1.data_ptr(), km + 1, 1, q.data_ptr(), km, 1, pe2.data_ptr(), kn + 1, 1, out.data_ptr(), kn, 1,
            ncol, km, kn, iv, kord]
    if name == "b200cs_mappm":
        args += [ws.data_ptr() if ws is not None else None, ws.numel() if ws is not None else 0]
    rc = getattr(lib, name)(*args, None)  # legacy default stream
    torch.cuda.synchronize()
    return rc, out


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from fv3net_b200 import _lib

    return _lib.load()


def test_mappm_static_matches_mappm_on_two_calls(lib):
    rng = np.random.default_rng(2026)
    # two calls in a row with different shapes and kord: the second finds the first one's leftovers in
    # the shared array (and a different [level][slot] layout), and must not care
    cases = [(3000, 79, 60, 9, 1), (1700, 40, 33, 7, 0), (3000, 79, 60, 12, -1), (500, 79, 79, 1, 1)]
    prepared = []
    for ncol, km, kn, kord, iv in cases:
        pe1, pe2 = _columns(rng, ncol, km, kn)
        q = (rng.normal(size=(ncol, km)) * 50).astype(np.float32)
        q[::5] = np.abs(q[::5]) * 1e-3
        dev = [torch.as_tensor(a).cuda() for a in (pe1, q, pe2)]
        nws = lib.b200cs_mappm_workspace_bytes(ncol, km) if kord > 6 else 0
        ws = torch.empty(max(nws, 1), dtype=torch.uint8, device="cuda")  # allocated HERE, before the calls under test
        rc, want = _call(lib, "b200cs_mappm", *dev, iv, kord, ws)
        assert rc == 0
        prepared.append((dev, iv, kord, want))
    before = torch.cuda.memory_allocated()
    for rep in range(2):
        for dev, iv, kord, want in prepared:
            rc, got = _call(lib, "b200cs_mappm_static", *dev, iv, kord)
            assert rc == 0, lib.b200cs_last_error()
            assert not torch.isnan(got).any()
            assert torch.equal(got, want), (rep, kord, iv)
            del got
    assert torch.cuda.memory_allocated() == before  # nothing of torch's grew either


def test_mappm_static_rejects_beyond_the_limit(lib):
    km, kn = 79, 10
    per_col = lib.b200cs_mappm_workspace_bytes(1, km)
    ncol = STATIC_MAX_BYTES // per_col + 1
    rng = np.random.default_rng(7)
    pe1, pe2 = _columns(rng, ncol, km, kn)
    q = rng.normal(size=(ncol, km)).astype(np.float32)
    dev = [torch.as_tensor(a).cuda() for a in (pe1, q, pe2)]
    rc, _ = _call(lib, "b200cs_mappm_static", *dev, 1, 9)
    assert rc == -4  # B200CS_EWORKSPACE
    msg = lib.b200cs_last_error().decode()
    assert "static array holds" in msg and str(STATIC_MAX_BYTES) in msg
    # exactly at the limit it runs, and kord <= 6 (no workspace) has no limit
    n_ok = STATIC_MAX_BYTES // per_col
    ws = torch.empty(lib.b200cs_mappm_workspace_bytes(n_ok, km), dtype=torch.uint8, device="cuda")
    sub = [a[:n_ok].contiguous() for a in dev]
    rc, want = _call(lib, "b200cs_mappm", *sub, 1, 9, ws)
    rc2, got = _call(lib, "b200cs_mappm_static", *sub, 1, 9)
    assert rc == 0 and rc2 == 0 and torch.equal(got, want)
    rc, want = _call(lib, "b200cs_mappm", *dev, 1, 4)
    rc2, got = _call(lib, "b200cs_mappm_static", *dev, 1, 4)
    assert rc == 0 and rc2 == 0 and torch.equal(got, want)
