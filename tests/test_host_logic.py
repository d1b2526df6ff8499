"""Host-side logic that needs no GPU: the C ABI library loads and exports every declared symbol,
parameter generation in the library matches the reference KATs, argument/shape errors are reported
as status codes, the device arithmetic's limb plumbing (host model) is exact, and the product
refuses to run without a CUDA device.  No compute calls."""
import ctypes as C
import os
import random
import re
import subprocess

import numpy as np
import pytest

from conftest import ROOT, load_golden, oracle_cfg, to_ints, to_limbs


@pytest.fixture(scope="module")
def S():
    import sponge_b200
    return sponge_b200


def test_library_exports_every_declared_symbol(S):
    hdr = open(os.path.join(ROOT, "include", "poseidon_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = sorted(set(re.findall(r"\b(psd_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 45
    lib = C.CDLL(S.LIB_PATH)
    missing = [n for n in declared if not hasattr(lib, n)]
    assert not missing, f"declared in include/poseidon_b200.h but not exported: {missing}"
    # and the Python binding prototypes cover the same set
    from sponge_b200 import _lib
    assert sorted(_lib._PROTOS) == declared


def test_version_and_status_strings(S):
    from sponge_b200._lib import lib
    assert b"sm_100a" in lib.psd_version()
    assert lib.psd_status_string(0) == b"PSD_OK" and lib.psd_status_string(3) == b"PSD_ERR_NONCANONICAL"


def test_library_parameter_generation_matches_reference_kats(S):
    """psd_find_poseidon_ark_and_mds / psd_default_params (C++ Grain LFSR) vs traits.rs:163-358."""
    entries = load_golden("reference_kats.json")["default_parameters"]["entries"]
    for e in entries:
        cfg = S.get_default_poseidon_parameters("bls12_381", e["rate"], e["optimized_for_weights"])
        assert str(cfg.ark[0][0]) == e["ark00"] and str(cfg.mds[0][0]) == e["mds00"]
        assert cfg.capacity == 1 and cfg.rate == e["rate"]
    assert S.get_default_poseidon_parameters("bls12_381", 9, False) is None
    assert S.get_default_poseidon_parameters("bls12_381", 1, False) is None


def test_library_parameter_generation_matches_oracle_fully(S, ref):
    for name, args in [("bn254_t3", ("bn254", 254, 2, 8, 57, 0)), ("bn254_t5", ("bn254", 254, 4, 8, 60, 0)),
                       ("bls_default_r3", ("bls12_381", 255, 3, 8, 56, 0))]:
        ark, mds = S.find_poseidon_ark_and_mds(*args)
        o = oracle_cfg(ref, name)
        assert ark == o.ark and mds == o.mds
    # skip_matrices consumes 2t samples per skipped matrix
    ark1, mds1 = S.find_poseidon_ark_and_mds("bn254", 254, 2, 8, 57, 1)
    oark, omds = ref.find_poseidon_ark_and_mds(ref.BN254_FR, 254, 2, 8, 57, 1)
    assert ark1 == oark and mds1 == omds
    with pytest.raises(S.PoseidonB200Error) as e:
        S.find_poseidon_ark_and_mds("bn254", 255, 2, 8, 57, 0)   # assert_eq!(MODULUS_BIT_SIZE, prime_num_bits)
    assert e.value.status == 2
    d = S.get_default_poseidon_parameters("bn254", 2, False)
    assert (d.full_rounds, d.partial_rounds, d.alpha) == (8, 57, 5) and d.ark == oracle_cfg(ref, "bn254_t3").ark


def test_config_shape_asserts(S):
    """PoseidonConfig::new asserts, mod.rs:196-203."""
    ok = dict(full_rounds=2, partial_rounds=1, alpha=5, mds=[[1, 2], [3, 4]], ark=[[1, 2]] * 3, rate=1, capacity=1)
    S.PoseidonConfig(**ok)
    for bad in (dict(ok, ark=[[1, 2]] * 2), dict(ok, ark=[[1]] * 3), dict(ok, mds=[[1, 2]]), dict(ok, mds=[[1], [2]])):
        with pytest.raises(AssertionError):
            S.PoseidonConfig(**bad)
    assert S.PoseidonParameters is S.PoseidonConfig


def test_product_fails_loudly_without_gpu(S, ref):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    cfg = S.get_default_poseidon_parameters("bls12_381", 2, False)
    with pytest.raises(S.PoseidonB200Error) as e:
        S.PoseidonSponge.new(cfg)
    assert e.value.status == 4 and "no CPU path" in str(e.value)


def test_c_abi_argument_errors(S):
    from sponge_b200 import _lib as L
    lib = L.lib
    h = C.c_void_p()
    assert lib.psd_ctx_create(None, 0, C.byref(h)) == L.PSD_ERR_ARG
    ark = np.zeros((3 * 3, 4), dtype=np.uint64)
    mds = np.zeros((9, 4), dtype=np.uint64)
    p = L.PsdParams(7, 2, 1, 5, 2, 1, ark.ctypes.data, mds.ctypes.data, 0)
    assert lib.psd_ctx_create(C.byref(p), 0, C.byref(h)) == L.PSD_ERR_ARG            # unknown field
    p = L.PsdParams(0, 2, 1, 5, 2, 1, None, mds.ctypes.data, 0)
    assert lib.psd_ctx_create(C.byref(p), 0, C.byref(h)) == L.PSD_ERR_SHAPE          # ark missing
    p = L.PsdParams(0, 2, 1, 5, 12, 1, ark.ctypes.data, mds.ctypes.data, 0)
    assert lib.psd_ctx_create(C.byref(p), 0, C.byref(h)) == L.PSD_ERR_UNSUPPORTED    # t = 13
    p = L.PsdParams(0, 2, 1, 5, 0, 3, ark.ctypes.data, mds.ctypes.data, 0)
    assert lib.psd_ctx_create(C.byref(p), 0, C.byref(h)) == L.PSD_ERR_SHAPE          # rate 0
    assert lib.psd_permute_batch(None, ark.ctypes.data, 1, 0, 1, 0, None) == L.PSD_ERR_ARG
    assert lib.psd_batch_new(None, 1, C.byref(h)) == L.PSD_ERR_ARG
    assert lib.psd_sponge_new(None, C.byref(h)) == L.PSD_ERR_ARG
    assert b"null" in lib.psd_last_error_string()
    lib.psd_ctx_destroy(None)
    lib.psd_batch_free(None)
    lib.psd_sponge_free(None)
    assert lib.psd_batch_size(None) == 0


def test_limb_helpers_and_enums(S):
    vals = [0, 1, (1 << 255) - 19, S.MODULUS["bn254"] - 1]
    assert S.limbs_to_ints(S.ints_to_limbs(vals)) == vals
    assert S.ints_to_limbs([1 + (2 << 64)]).tolist() == [[1, 2, 0, 0]]
    assert S.FieldElementSize.Truncated(100).num_bits("bls12_381") == 254     # quirk Q3, lib.rs:45-52
    assert S.FieldElementSize.Full().num_bits("bn254") == 253
    with pytest.raises(ValueError):
        S.FieldElementSize.Truncated(256).num_bits("bls12_381")
    assert S.DuplexSpongeMode.Absorbing(0) != S.DuplexSpongeMode.Squeezing(0)


@pytest.fixture(scope="module")
def emu():
    # PSD_SANITIZE=1 (tests/test_sanitizers.py): the same host model built with ASan + UBSan
    san = os.environ.get("PSD_SANITIZE") == "1"
    so = os.path.join(ROOT, "tests", "emu", "libfp_emu_san.so" if san else "libfp_emu.so")
    csrc = os.path.join(ROOT, "sponge_b200", "csrc")
    srcs = [os.path.join(ROOT, "tests", "emu", "fp_emu.cpp"), os.path.join(csrc, "opt_constants.cpp"),
            os.path.join(csrc, "fp52_tables.cpp"), os.path.join(csrc, "tc_tables.cpp"),
            os.path.join(csrc, "params.cpp")]
    deps = srcs + [os.path.join(csrc, h) for h in ("fp256.cuh", "poseidon_core.cuh", "opt_constants.hpp",
                                                   "host_field.hpp", "params.hpp", "fp52.cuh", "poseidon_fp52.cuh",
                                                   "fp52_tables.hpp", "poseidon_tc.cuh", "tc_tables.hpp")]
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(d) for d in deps):
        # -frounding-math: fp52.cuh's host model runs fma() under FE_TOWARDZERO
        flags = ["-O1", "-g", "-fsanitize=address,undefined", "-fno-sanitize-recover=all"] if san else ["-O2"]
        subprocess.check_call(["g++"] + flags + ["-std=c++17", "-frounding-math", "-ffp-contract=off", "-shared", "-fPIC",
                               "-o", so] + srcs)
    return _PtrSafe(C.CDLL(so))


class _PtrSafe:
    """ctypes passes a bare Python int as a C int; buffer addresses (`arr.ctypes.data`) must go as pointers."""

    def __init__(self, lib):
        self._lib = lib

    def __getattr__(self, name):
        fn = getattr(self._lib, name)

        def call(*args):
            return fn(*[C.c_void_p(a) if isinstance(a, int) and not isinstance(a, bool) and a >= (1 << 31) else a
                        for a in args])
        return call


@pytest.mark.parametrize("field", [0, 1])
def test_device_arithmetic_host_model(emu, ref, field):
    """sponge_b200/csrc/fp256.cuh compiled for the host (C model of each PTX carry chain):
    Montgomery product, modular add and the pow ladder are exact, including edge operands."""
    p = ref.BN254_FR if field == 0 else ref.BLS12_381_FR
    rng = random.Random(field)
    n = 4000
    A = [rng.randrange(p) for _ in range(n)]
    B = [rng.randrange(p) for _ in range(n)]
    A[:8] = [0, p - 1, p - 1, 1, 0, p - 2, (1 << 253) - 1, 1 << 253]
    B[:8] = [0, p - 1, 1, p - 1, p - 1, p - 1, p - 1, (1 << 253) + 5]
    a, b = to_limbs(A).view(np.uint32), to_limbs(B).view(np.uint32)
    r = np.zeros_like(a)
    rinv = pow(1 << 256, -1, p)
    emu.emu_mont_mul(field, r.ctypes.data, a.ctypes.data, b.ctypes.data, C.c_size_t(n))
    assert to_ints(r.view(np.uint64)) == [x * y * rinv % p for x, y in zip(A, B)]
    emu.emu_add(field, r.ctypes.data, a.ctypes.data, b.ctypes.data, C.c_size_t(n))
    assert to_ints(r.view(np.uint64)) == [(x + y) % p for x, y in zip(A, B)]
    one = to_limbs([(1 << 256) % p]).view(np.uint32)
    for alpha in (5, 17, 257, 3, 1, 0):
        emu.emu_pow(field, r.ctypes.data, a.ctypes.data, C.c_uint64(alpha), one.ctypes.data, C.c_size_t(400))
        exp = [pow(x * rinv % p, alpha, p) * (1 << 256) % p for x in A[:400]]
        assert to_ints(r.view(np.uint64))[:400] == exp


@pytest.mark.parametrize("name", ["bn254_t3", "bn254_t5", "bls_default_r2", "bls_fixture", "bn254_t2",
                                  "bn254_t4_cap2", "bn254_odd_rf", "bls_default_r2_w", "bn254_full_only",
                                  "bn254_rf1", "bls_t3_alpha5", "bn254_rp1", "bn254_rp2_alpha17"])
def test_permutation_variants_host_model(emu, ref, name):
    """The device permutation code (poseidon_core.cuh: dense, optimised sparse/lazy form, rolled form)
    compiled for the host with the product's own constant derivation (opt_constants.cpp), against the
    oracle.  This checks the algebra of the optimised form without a GPU; the PTX mapping itself is
    covered by the -m gpu tests."""
    from conftest import field_id
    from oracle.c_oracle import COracle
    cfg = oracle_cfg(ref, name)
    fid = field_id(ref, cfg)
    t = cfg.rate + cfg.capacity
    orc = COracle(cfg, fid)
    n = 24
    st = orc.synth(n * t, seed=5)
    st[:t] = 0
    st[t:2 * t] = to_limbs([cfg.p - 1] * t)      # largest residues in every lane
    exp = orc.permute_batch(st.copy())
    ark = to_limbs(v for r in cfg.ark for v in r)
    mds = to_limbs(v for r in cfg.mds for v in r)
    variants = [0, 1, 2, 3]          # dense, optimised, rolled, compact optimised
    if t <= 3:
        variants.append(4)           # compact optimised with digit-table constant products
    if cfg.partial_rounds and cfg.full_rounds >= 2 and t <= 5:
        variants.append(5)           # ... with tables for the lane updates only (the t = 5 mode)
    if cfg.alpha in (5, 17) and t <= 5 and cfg.full_rounds >= 2:   # t = 5: the shared-memory form's tables
        variants.append(6)           # linear layer in 52-bit limbs on the FP64 pipe (fma under round-toward-zero)
    for variant in variants:
        s = st.copy().view(np.uint32)
        rc = emu.emu_permute(fid, cfg.rate, cfg.capacity, cfg.full_rounds, cfg.partial_rounds, C.c_uint64(cfg.alpha),
                             C.c_void_p(ark.ctypes.data), C.c_void_p(mds.ctypes.data), variant,
                             C.c_void_p(s.ctypes.data), C.c_size_t(n))
        if rc == 1:      # the requested optimised form does not exist for this parameter set (e.g. full_rounds < 2)
            assert cfg.full_rounds < 2
            continue
        assert rc == 0, (variant, rc)
        assert np.array_equal(s.view(np.uint64), exp), f"variant {variant} differs from the oracle"


def test_lazy_sbox_host_model(emu, ref):
    """BN254 S-box with the inner conditional subtractions dropped (digit-table variant): the raw result is
    congruent to x^alpha, stays below the 1.26 p bound stated in poseidon_core.cuh, also for x = p - 1."""
    p = ref.BN254_FR
    rng = random.Random(77)
    rinv = pow(1 << 256, -1, p)
    n = 3000
    A = [rng.randrange(p) for _ in range(n)]
    A[:8] = [0, 1, p - 1, p - 2, (1 << 253) - 1, 1 << 253, 0xffffffff, p // 2]
    a = to_limbs(A).view(np.uint32)
    for alpha in (5, 17):
        r = np.zeros_like(a)
        emu.emu_sbox_lazy_bn254(C.c_uint64(alpha), C.c_void_p(r.ctypes.data), C.c_void_p(a.ctypes.data), C.c_size_t(n))
        got = to_ints(r.view(np.uint64))
        for x, g in zip(A, got):
            assert g % p == pow(x, alpha, p) * pow(rinv, alpha - 1, p) % p
            assert g * 100 < 126 * p


def test_mont_dot_host_model(emu, ref):
    """Lazily reduced dot product (one Montgomery reduction for N products), BN254, worst-case operands."""
    p = ref.BN254_FR
    rng = random.Random(4)
    rinv = pow(1 << 256, -1, p)
    for n_terms in (1, 2, 3, 4):
        n = 1500
        A = [[rng.randrange(p) for _ in range(n_terms)] for _ in range(n)]
        B = [[rng.randrange(p) for _ in range(n_terms)] for _ in range(n)]
        A[0] = [p - 1] * n_terms
        B[0] = [p - 1] * n_terms
        A[1] = [0] * n_terms
        a = to_limbs(v for r in A for v in r).view(np.uint32)
        b = to_limbs(v for r in B for v in r).view(np.uint32)
        r = np.zeros(n * 8, dtype=np.uint32)
        emu.emu_mont_dot(n_terms, C.c_void_p(r.ctypes.data), C.c_void_p(a.ctypes.data), C.c_void_p(b.ctypes.data),
                         C.c_size_t(n))
        assert to_ints(r.view(np.uint64)) == [sum(x * y for x, y in zip(ra, rb)) * rinv % p for ra, rb in zip(A, B)]


def _build_cpp_wrapper_test(S):
    exe = os.path.join(ROOT, "tests", "cpp", "wrapper_test")
    src = exe + ".cpp"
    hdrs = [os.path.join(ROOT, "include", h) for h in ("poseidon_b200.hpp", "poseidon_b200.h")]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(p) for p in [src, S.LIB_PATH] + hdrs):
        libdir = os.path.dirname(S.LIB_PATH)
        subprocess.check_call(["g++", "-std=c++17", "-Wall", "-I" + os.path.join(ROOT, "include"), "-o", exe, src,
                               "-L" + libdir, "-lposeidon_b200", "-Wl,-rpath," + libdir])
    return exe


def test_cpp_wrapper_compiles_and_refuses_without_gpu(S):
    """include/poseidon_b200.hpp builds against the C ABI; without a GPU the program reports PSD_ERR_CUDA
    (exit code 3) instead of computing anything on the CPU."""
    import torch
    exe = _build_cpp_wrapper_test(S)
    kat = load_golden("reference_kats.json")["sponge_consistency"]["squeeze_native"]
    r = subprocess.run([exe] + ["%064x" % int(v) for v in kat], capture_output=True, text=True)
    if torch.cuda.is_available():
        assert r.returncode == 0, r.stdout
    else:
        assert r.returncode == 3 and "no CPU path" in r.stdout


@pytest.mark.parametrize("field", [0, 1])
def test_squaring_and_wide_window_host_model(emu, ref, field):
    """Dedicated squaring (36 wide multiplies) and the 10-limb-window variants used for BLS12-381
    (lazy dot products whose running sum exceeds 9 limbs, results up to 3p), worst-case operands."""
    p = ref.BN254_FR if field == 0 else ref.BLS12_381_FR
    rng = random.Random(100 + field)
    rinv = pow(1 << 256, -1, p)
    n = 4000
    A = [rng.randrange(p) for _ in range(n)]
    A[:8] = [0, 1, p - 1, p - 2, (1 << 253) - 1, 1 << 253, 0xffffffff, 1 << 32]
    a = to_limbs(A).view(np.uint32)
    exp = [x * x * rinv % p for x in A]
    for fn in (emu.emu_mont_sqr, emu.emu_mont_sqr_w):
        r = np.zeros_like(a)
        fn(field, C.c_void_p(r.ctypes.data), C.c_void_p(a.ctypes.data), C.c_size_t(n))
        assert to_ints(r.view(np.uint64)) == exp
    for n_terms in (1, 2, 3, 4):
        m = 1500
        AA = [[rng.randrange(p) for _ in range(n_terms)] for _ in range(m)]
        BB = [[rng.randrange(p) for _ in range(n_terms)] for _ in range(m)]
        AA[0] = [p - 1] * n_terms
        BB[0] = [p - 1] * n_terms
        AA[1] = [0] * n_terms
        av = to_limbs(v for r_ in AA for v in r_).view(np.uint32)
        bv = to_limbs(v for r_ in BB for v in r_).view(np.uint32)
        r = np.zeros(m * 8, dtype=np.uint32)
        emu.emu_mont_dot_w(field, n_terms, C.c_void_p(r.ctypes.data), C.c_void_p(av.ctypes.data),
                           C.c_void_p(bv.ctypes.data), C.c_size_t(m))
        assert to_ints(r.view(np.uint64)) == [sum(x * y for x, y in zip(ra, rb)) * rinv % p for ra, rb in zip(AA, BB)]


def test_fp52_primitives_host_model(emu):
    """fp52.cuh: the two-FMA exact product (round toward zero) and the 52-bit limb packing, on extreme and
    random operands, against Python integers."""
    rng = random.Random(52)
    top = (1 << 52) - 1
    edge = [0, 1, 2, top, top - 1, 1 << 51, (1 << 51) - 1, (1 << 26) - 1, 1 << 26, 0xfffff00000000, 0xfffffffff]
    A = [x for x in edge for _ in edge] + [rng.randrange(1 << 52) for _ in range(20000)]
    B = [y for _ in edge for y in edge] + [rng.randrange(1 << 52) for _ in range(20000)]
    n = len(A)
    a = np.array(A, dtype=np.uint64)
    b = np.array(B, dtype=np.uint64)
    hi = np.zeros(n, dtype=np.uint64)
    lo = np.zeros(n, dtype=np.uint64)
    emu.emu_f52_prod(C.c_void_p(a.ctypes.data), C.c_void_p(b.ctypes.data), C.c_void_p(hi.ctypes.data),
                     C.c_void_p(lo.ctypes.data), C.c_size_t(n))
    for x, y, h, l in zip(A, B, hi.tolist(), lo.tolist()):
        assert (h << 52) + l == x * y and l < (1 << 52)
    vals = [0, 1, (1 << 256) - 1, (1 << 52) - 1, 1 << 52, (1 << 104) - 1, 1 << 208, (1 << 208) - 1] + \
           [rng.randrange(1 << 256) for _ in range(2000)]
    x = to_limbs(vals).view(np.uint32)
    limbs = np.zeros(5 * len(vals), dtype=np.uint64)
    back = np.zeros_like(x)
    emu.emu_f52_unpack_pack(C.c_void_p(x.ctypes.data), C.c_void_p(limbs.ctypes.data), C.c_void_p(back.ctypes.data),
                            C.c_size_t(len(vals)))
    assert np.array_equal(back, x)
    lm = limbs.reshape(-1, 5).tolist()
    for v, row in zip(vals, lm):
        assert all(r < (1 << 52) for r in row[:4]) and sum(r << (52 * k) for k, r in enumerate(row)) == v



@pytest.mark.parametrize("name", ["bn254_t3", "bn254_t5", "bn254_t2", "bn254_rp2_alpha17", "bn254_rp1", "bn254_full_only"])
def test_fp52_bounds_are_evaluated_on_the_tables(emu, ref, name):
    """Variant 5 is exact only under numeric bounds (table limbs < 2^52, entries < p, rows < p + 2^207, lazily reduced
    lanes < 2^258, no column wrap, lazy S-box range).  The library evaluates them on the image it is about to upload
    (fp52_image_bounds_ok, called from append_fp52_tables; a failing image leaves fp_ok = 0 and the context runs an
    integer variant).  Here: the real tables pass; damaged ones -- a limb of 2^52, an all-ones table, a wrong modulus
    limb, an additive constant past 2^52 -- are refused with the bound named; lane-update tables replaced by the
    largest residue p - 1 still pass (the bounds hold for ANY residues)."""
    cfg = oracle_cfg(ref, name)
    ark = to_limbs(v for r in cfg.ark for v in r)
    mds = to_limbs(v for r in cfg.mds for v in r)
    why = C.create_string_buffer(200)

    def run(corrupt):
        rc = emu.emu_fp52_bounds(0, cfg.rate, cfg.capacity, cfg.full_rounds, cfg.partial_rounds, C.c_void_p(ark.ctypes.data),
                                 C.c_void_p(mds.ctypes.data), corrupt, why, C.c_size_t(200))
        return rc, why.value.decode()
    assert run(0) == (1, "")
    assert run(4)[0] == 1
    for corrupt, word in ((1, "MDS tables"), (2, "MDS tables"), (3, "modulus"), (5, "additive constant")):
        rc, msg = run(corrupt)
        assert rc == 0 and word in msg, (corrupt, rc, msg)


@pytest.mark.parametrize("name", ["bn254_t3", "bn254_t5", "bn254_t2", "bn254_t4_cap2", "bn254_rp2_alpha17", "bn254_rp1",
                                  "bn254_full_only", "bls_fixture", "bls_default_r2", "bls_t3_alpha5", "bn254_odd_rf"])
def test_tensor_core_variant_host_model(emu, ref, name):
    """Variant 6 without a GPU: the product's own byte-digit tables (tc_tables.cpp, in mma fragment order) driven per
    sponge by the host model of one output value -- column sums, carry sweep, one 32-bit Montgomery row -- and the
    device's integer S-box code, against the oracle.  Checks the table construction, the ordering of the additive
    constants and the bounds; the fragment plumbing (mma.sync, ldmatrix, shared-memory exchange) is covered by the
    -m gpu tests."""
    from conftest import field_id
    from oracle.c_oracle import COracle
    cfg = oracle_cfg(ref, name)
    fid = field_id(ref, cfg)
    t = cfg.rate + cfg.capacity
    orc = COracle(cfg, fid)
    n = 24
    st = orc.synth(n * t, seed=6)
    st[:t] = 0
    st[t:2 * t] = to_limbs([cfg.p - 1] * t)
    exp = orc.permute_batch(st.copy())
    ark = to_limbs(v for r in cfg.ark for v in r)
    mds = to_limbs(v for r in cfg.mds for v in r)
    for layout in (0, 1):   # mma.sync fragment order (variant 6) and tcgen05 core-matrix blocks (variant 7)
        s = st.copy().view(np.uint32)
        rc = emu.emu_permute_tc_layout(fid, cfg.rate, cfg.capacity, cfg.full_rounds, cfg.partial_rounds, C.c_uint64(cfg.alpha),
                                       C.c_void_p(ark.ctypes.data), C.c_void_p(mds.ctypes.data), C.c_void_p(s.ctypes.data),
                                       C.c_size_t(n), layout)
        if cfg.alpha not in (5, 17):
            assert rc == 1
            continue
        assert rc == 0
        assert np.array_equal(s.view(np.uint64), exp), layout


def test_unreduced_products(emu):
    """mul_wide / sqr_wide (fp256.cuh, the S-box of variant 7): exact 512-bit products, worst-case limbs included."""
    rng = np.random.default_rng(11)
    n = 64
    a = rng.integers(0, 2**32, size=(n, 8), dtype=np.uint64).astype(np.uint32)
    b = rng.integers(0, 2**32, size=(n, 8), dtype=np.uint64).astype(np.uint32)
    a[0] = 0xffffffff
    b[0] = 0xffffffff
    a[1] = 0
    a[2] = 0xffffffff
    a[2, 7] = 0x7fffffff
    a[3] = 0x80000000
    a[3, 7] = 0x40000000

    def ints(x):
        return [sum(int(w) << (32 * k) for k, w in enumerate(row)) for row in x]
    out = np.zeros((n, 16), dtype=np.uint32)
    emu.emu_mul_wide(C.c_void_p(a.ctypes.data), C.c_void_p(b.ctypes.data), C.c_void_p(out.ctypes.data), C.c_size_t(n))
    assert ints(out) == [x * y for x, y in zip(ints(a), ints(b))]
    sq_in = a.copy()
    sq_in[:, 7] &= 0x7fffffff   # the squaring's operand range: below 2^255
    emu.emu_mul_wide(C.c_void_p(sq_in.ctypes.data), None, C.c_void_p(out.ctypes.data), C.c_size_t(n))
    assert ints(out) == [x * x for x in ints(sq_in)]
