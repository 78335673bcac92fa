"""The exact replay of the rolling window energy (walk.cu: exact_energy = lane-parallel rounds + literal rounds)
against the oracle's literal chain (decoder.rs:133-136, 224-233), bit for bit, on captures built to stress the
translation-invariance argument of the lane-parallel rounds: energies sitting on powers of two, amplitude steps
of many binades, silence (clamp / gate), long chains, NaN / inf, misaligned seeds."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
MAN, FB = 0, 1


def probe(engine, orc, cfg, x, seed, windows, ctx):
    import tmkphy
    P = 60 if cfg.line_code == FB else 96
    assert P == int(tmkphy._abi.load_library().tmk_preamble_len(cfg))
    got, cnt = engine.replay_probe(cfg, x, seed, windows)
    want = orc.rolling_energy(x, P, seed, windows)
    same = (got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want))
    assert same.all(), f"{ctx}: E_ref differs at window {np.asarray(windows)[~same][0]}: {got[~same][0]!r} vs {want[~same][0]!r}"
    return cnt


def noise(rng, n, sigma):
    return rng.normal(0.0, sigma, n).astype(np.float32)


@pytest.mark.parametrize("kind", [FB, MAN])
@pytest.mark.parametrize("seed_off", [0, 1, 2, 3, 1201])
def test_replay_long_noise_chain(engine, orc, kind, seed_off):
    """300 k steps of noise: the lane-parallel rounds carry almost all of it"""
    import tmkphy
    cfg = tmkphy.make_config(kind)
    rng = np.random.default_rng(10 + seed_off)
    x = noise(rng, 300_000, 0.37)
    ws = np.array([seed_off + 5, 4200 + seed_off, 70_001, 70_002, 150_000, 299_000], np.uint64)
    cnt = probe(engine, orc, cfg, x, seed_off, ws, f"noise/{kind}/{seed_off}")
    assert cnt["parallel_steps"] > 200_000


@pytest.mark.parametrize("target", [1.0, 2.0, 32.0, 0.5, 2.0 ** -10])
def test_replay_energy_on_a_power_of_two(engine, orc, target):
    """window energy hovering exactly around a binade boundary: e - a falls below it, + b rises above it at
    every step, so most segments must be recognised as in doubt (and still come out exact)"""
    import tmkphy
    cfg = tmkphy.make_config(FB)
    rng = np.random.default_rng(int(target * 1024) + 5)
    sigma = float(np.sqrt(target / 60.0))
    x = noise(rng, 200_000, sigma)
    ws = np.arange(5_000, 199_000, 9_973, dtype=np.uint64)
    cnt = probe(engine, orc, cfg, x, 0, ws, f"pow2/{target}")
    assert cnt["parallel_steps"] > 100_000 and cnt["parallel_doubts"] > 0


def test_replay_amplitude_steps_silence_and_bursts(engine, orc):
    """binade jumps inside segments, digital silence (the clamp and E == 0), single loud samples"""
    import tmkphy
    cfg = tmkphy.make_config(FB)
    rng = np.random.default_rng(77)
    parts = [noise(rng, 40_000, 0.3), noise(rng, 30_000, 3e-3), np.zeros(20_000, np.float32), noise(rng, 40_000, 0.3),
             noise(rng, 30_000, 40.0), noise(rng, 40_000, 1e-4), noise(rng, 50_000, 0.7)]
    x = np.concatenate(parts)
    x[55_555] = 900.0       # a burst: the window energy jumps by 20 binades and comes back
    x[200_001] = -3.0e4
    x[120_000:120_400] = 0.0
    ws = np.arange(1_000, len(x) - 100, 4_999, dtype=np.uint64)
    probe(engine, orc, cfg, x, 0, ws, "steps")
    probe(engine, orc, cfg, x, 39_990, ws[ws > 40_000], "steps, seeded before the amplitude step")


def test_replay_nan_inf_and_tiny(engine, orc):
    import tmkphy
    cfg = tmkphy.make_config(MAN)
    rng = np.random.default_rng(3)
    x = noise(rng, 150_000, 0.25)
    ws = np.arange(500, 149_000, 7_001, dtype=np.uint64)
    y = x.copy()
    y[80_000] = np.inf
    probe(engine, orc, cfg, y, 0, ws, "inf")
    y = x.copy()
    y[30_003] = np.nan
    probe(engine, orc, cfg, y, 0, ws, "nan")
    probe(engine, orc, cfg, (x * np.float32(1e-18)).astype(np.float32), 0, ws, "subnormal squares")
    probe(engine, orc, cfg, (x * np.float32(1e17)).astype(np.float32), 0, ws, "huge")


def test_replay_resume_and_short_chains(engine, orc):
    """many closely spaced queries (the cached chain position is resumed), chains below the parallel threshold"""
    import tmkphy
    cfg = tmkphy.make_config(FB)
    rng = np.random.default_rng(9)
    x = noise(rng, 60_000, 0.5)
    ws = np.sort(rng.integers(100, 59_000, 400)).astype(np.uint64)
    probe(engine, orc, cfg, x, 100, ws, "resume")
    probe(engine, orc, cfg, x, 7, np.array([7, 8, 100, 4000, 4103, 4104, 4105], np.uint64), "short")
