"""Worst-case intermediate states for the shipped fast kernels (variants 6 and 5).

Random inputs almost never put an extreme value INSIDE the permutation: the state after the S-box of round r
is uniformly random whatever the input was.  Here the extreme values are placed there on purpose: a target
state right after the S-box layer of a chosen round (lane values 0, p-1, 2^253, values whose Montgomery image
is p-1 / 2^253 / 2^208-1 / all 52-bit limbs set, ...) is mapped BACK to the permutation's input by running the
oracle's rounds in reverse -- inverse MDS, minus the round constants, alpha-th root (alpha^-1 mod p-1) -- so
that the forward run passes through exactly that state.  Rounds: the first and last of both full-round groups
and the first, second, a middle and the last two partial rounds.  The GPU result (variants 6 and 5, and variant 3 as
the all-integer cross-check) must equal the Python oracle's forward permutation bit for bit, and the oracle's
own forward run is asserted to hit the target, so the construction itself is checked.

Reference being restated by the kernel: /root/reference/src/poseidon/mod.rs:63-118."""
import numpy as np
import pytest

from conftest import field_name, oracle_cfg, to_ints, to_limbs

pytestmark = pytest.mark.gpu


def mat_inv(m, p):
    n = len(m)
    a = [list(r) + [1 if i == j else 0 for j in range(n)] for i, r in enumerate(m)]
    for c in range(n):
        piv = next(r for r in range(c, n) if a[r][c] % p)
        a[c], a[piv] = a[piv], a[c]
        inv = pow(a[c][c], -1, p)
        a[c] = [x * inv % p for x in a[c]]
        for r in range(n):
            if r != c and a[r][c]:
                f = a[r][c]
                a[r] = [(x - f * y) % p for x, y in zip(a[r], a[c])]
    return [r[n:] for r in a]


def input_for_state_after_sbox(cfg, r_target, y):
    """Input x such that the oracle's state right after apply_s_box of round r_target equals y."""
    p, t = cfg.p, cfg.rate + cfg.capacity
    half = cfg.full_rounds // 2
    inv_alpha = pow(cfg.alpha, -1, p - 1)
    minv = mat_inv(cfg.mds, p)
    st = list(y)
    for r in range(r_target, -1, -1):
        full = r < half or r >= half + cfg.partial_rounds
        if full:
            st = [pow(v, inv_alpha, p) for v in st]
        else:
            st[0] = pow(st[0], inv_alpha, p)
        st = [(st[i] - cfg.ark[r][i]) % p for i in range(t)]
        if r > 0:
            st = [sum(minv[i][j] * st[j] for j in range(t)) % p for i in range(t)]
    return st


def state_after_sbox(cfg, x, r_target):
    p, t = cfg.p, cfg.rate + cfg.capacity
    half = cfg.full_rounds // 2
    st = list(x)
    for r in range(r_target + 1):
        full = r < half or r >= half + cfg.partial_rounds
        st = [(st[i] + cfg.ark[r][i]) % p for i in range(t)]
        if full:
            st = [pow(v, cfg.alpha, p) for v in st]
        else:
            st[0] = pow(st[0], cfg.alpha, p)
        if r == r_target:
            return st
        st = [sum(st[j] * cfg.mds[i][j] for j in range(t)) % p for i in range(t)]


@pytest.mark.parametrize("name", ["bn254_t3", "bn254_t5", "bn254_t2", "bn254_rp2_alpha17"])
def test_extreme_values_inside_the_permutation(ref, name):
    import sponge_b200 as S
    cfg = oracle_cfg(ref, name)
    p, t = cfg.p, cfg.rate + cfg.capacity
    half, rp, total = cfg.full_rounds // 2, cfg.partial_rounds, cfg.full_rounds + cfg.partial_rounds
    rinv = pow(1 << 256, -1, p)
    # true values; "img" ones are chosen by their Montgomery image (what the kernel's registers hold)
    plain = [0, 1, p - 1, p - 2, 1 << 253, (1 << 253) - 1, (p - 1) // 2, (p + 1) // 2]
    images = [p - 1, 1 << 253, (1 << 208) - 1, (1 << 52) - 1, 1 << 52, (1 << 104) - 1, (1 << 156) - 1,
              ((1 << 253) - 1) & ~((1 << 52) - 1), sum(((1 << 52) - 1) << (52 * k) for k in range(4)) + ((1 << 45) << 208)]
    values = plain + [v % p * rinv % p for v in images]
    rounds = sorted({0, half - 1, half, half + 1, half + rp // 2, half + rp - 2, half + rp - 1, half + rp, total - 1}
                    & set(range(total)))
    inputs, targets = [], []
    for r in rounds:
        for k, v in enumerate(values):
            # the extreme value in every lane; in one lane only (lane 0 = the partial rounds' S-box lane); mixed
            for y in ([v] * t, [v] + [values[(k + 3 + j) % len(values)] for j in range(1, t)],
                      [values[(k + 5) % len(values)]] * (t - 1) + [v]):
                x = input_for_state_after_sbox(cfg, r, y)
                inputs.append(x)
                targets.append((r, y))
    # the construction really goes through the target state (checked on a sample; every 7th)
    for x, (r, y) in list(zip(inputs, targets))[::7]:
        assert state_after_sbox(cfg, x, r) == y
    expected = [ref.permute(cfg, x) for x in inputs]
    flat_in = to_limbs(v for x in inputs for v in x)
    pcfg = S.PoseidonConfig(cfg.full_rounds, cfg.partial_rounds, cfg.alpha, cfg.mds, cfg.ark, cfg.rate, cfg.capacity,
                            field_name(ref, cfg))
    ctx = pcfg.context()
    try:
        for variant in (7, 6, 5, 3, 0):
            ctx.set_variant(variant)
            if variant >= 5:
                assert ctx.active_variant() == variant, ctx.describe()
            got = to_ints(ctx.permute_batch(flat_in.copy(), repr=S.REPR_CANONICAL))
            got = [got[i * t:(i + 1) * t] for i in range(len(inputs))]
            bad = [i for i in range(len(inputs)) if got[i] != expected[i]]
            assert not bad, f"variant {variant}: {len(bad)} of {len(inputs)} differ, first: round {targets[bad[0]][0]}"
    finally:
        ctx.close()
