"""CPU: host-side logic that needs no GPU — wire codec, the contract mirror's activation gate and vote
state machine, trial sharding and the world_size-2 all-reduce of the accumulators (gloo)."""
import os
import socket

import numpy as np
import pytest
import torch


def test_felt_codec_round_trips():
    from svoc import client
    P = client.UPPER_BOUND_FELT252
    assert client.felt_to_i128(P - 1) == -1 and client.i128_to_felt(-1) == P - 1          # tests/test_math.cairo:95-96
    assert client.float_to_fwsad(0.492954) == 492954 and client.float_to_fwsad(-2.5) == P - 2500000
    assert client.fwsad_to_float(client.float_to_fwsad(-2.5)) == -2.5
    assert client.float_to_fwsad(0.9999999) == 999999                                      # int() truncates toward zero
    assert client.from_hex(client.to_hex(255)) == 255
    for v in (0, 1, -1, 10**12, -(10**12), 2**126):
        assert client.felt_to_i128(client.i128_to_felt(v)) == v


def test_contract_mirror_gate_and_vote_without_gpu():
    from svoc.contract import ContractPanic, OracleConsensusNDS
    admins, oracles = ["Akashi", "Ozu", "Higuchi"], [f"oracle_{i:02d}" for i in range(7)]
    c = OracleConsensusNDS(admins, True, 2, 2, True, 0, 2, oracles)
    for i in range(6):                                  # no consensus (and no GPU call) until all have committed
        c.set_caller(oracles[i])
        c.update_prediction([400000 + i, 300000])
        assert not c.consensus_active() and c.get_consensus_value() == [0, 0]
    assert c.n_active_oracles == 6
    c.set_caller(oracles[0])
    c.update_prediction([1, 2])                          # re-commit does not count twice
    assert c.n_active_oracles == 6
    with pytest.raises(ContractPanic):
        c.update_prediction([-1 % 2**251, 5])            # negative felt -> interval error
    c.set_caller("stranger")
    with pytest.raises(ContractPanic):
        c.update_prediction([1, 2])
    with pytest.raises(ContractPanic):
        c.update_proposition((1, "x"))                   # not an admin
    with pytest.raises(ContractPanic):
        c.get_oracle_value_list()
    c.set_caller(admins[2])
    with pytest.raises(ContractPanic):
        c.update_proposition((7, "x"))                   # wrong old oracle index
    with pytest.raises(ContractPanic):
        c.update_proposition((1, oracles[3]))            # already in the team
    c.update_proposition((1, "new_1"))
    assert c.get_a_specific_proposition(2) == (1, "new_1") and c.vote_matrix[2][2]
    c.update_proposition((2, "new_2"))                   # changing the proposition resets its votes
    c.set_caller(admins[0])
    c.vote_for_a_proposition(2, True)
    assert c.get_oracle_list()[2] == "new_2" and c.get_replacement_propositions() == [None] * 3
    assert not any(any(r) for r in c.vote_matrix)
    c.vote_for_a_proposition(1, True)                    # 1 vote < majority: nothing happens
    c.set_caller(admins[1])
    with pytest.raises(ContractPanic):
        c.vote_for_a_proposition(1, True)                # majority reached on an empty proposition: unwrap panics
    assert not c.vote_matrix[1][1] and c.vote_matrix[0][1]   # the reverted transaction left no vote behind
    c.vote_for_a_proposition(7, True)                    # a receiver beyond the admin list: a storage map takes any key
    c.set_caller(admins[2])
    with pytest.raises(ContractPanic):
        c.vote_for_a_proposition(7, True)                # second vote at that key: majority on a proposition that is None
    assert c._stray_votes == {(1, 7): True, (2, 7): False}
    d = OracleConsensusNDS(admins, False, 2, 2, True, 0, 2, oracles)
    d.set_caller(admins[0])
    with pytest.raises(ContractPanic):
        d.update_proposition(None)                       # replacement disabled


def test_shard_ranges_partition_the_trials():
    from svoc.distributed import shard_range
    for T in (0, 1, 7, 1000, 10**9 + 7):
        for G in (1, 2, 3, 8):
            r = [shard_range(T, k, G) for k in range(G)]
            assert r[0][0] == 0 and r[-1][1] == T
            assert all(r[k][1] == r[k + 1][0] for k in range(G - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, T, out):
    import torch.distributed as dist
    from oracle import cref
    from svoc.distributed import allreduce_stats, shard_range
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(42)
        x = (rng.beta(10, 15, (T, 7, 2)) * 1e6).astype(np.int64)        # same data on every rank, sharded by range
        lo, hi = shard_range(T, rank, world)
        o = cref.fixed_batch(x[lo:hi], 2, True, 0, False)
        counts = torch.zeros(528 + 14, dtype=torch.int64)
        sums = torch.zeros(4 + 7, dtype=torch.float64)
        ok = o["status"] == 0
        counts[0], counts[1] = hi - lo, int(ok.sum())
        counts[528:535] = torch.from_numpy((~o["reliable"][ok]).sum(axis=0))
        counts[535:542] = torch.from_numpy(o["qr"][ok].sum(axis=0))
        sums[0] = float(o["rel1"][ok].sum()) * 1e-6
        counts[540] = (1 << 60) + 12345 + rank        # beyond 2^53: must survive the float64 transport exactly
        c, s = allreduce_stats(counts, sums)[:2]
        if rank == 0:
            torch.save({"counts": c, "sums": s}, out)
    finally:
        dist.destroy_process_group()


def test_world_size_2_allreduce_equals_single_rank(tmp_path):
    """Two gloo ranks each reduce their own trial range; the all-reduced accumulators equal the
    single-process result exactly (integer counters) / to rounding (float sums)."""
    import torch.multiprocessing as mp
    from oracle import cref
    T, port, out = 1001, _free_port(), str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, port, T, out), nprocs=2, join=True)
    got = torch.load(out)
    rng = np.random.default_rng(42)
    x = (rng.beta(10, 15, (T, 7, 2)) * 1e6).astype(np.int64)
    o = cref.fixed_batch(x, 2, True, 0, False)
    ok = o["status"] == 0
    assert int(got["counts"][0]) == T and int(got["counts"][1]) == int(ok.sum())
    assert got["counts"][528:535].tolist() == (~o["reliable"][ok]).sum(axis=0).tolist()
    want = o["qr"][ok].sum(axis=0).tolist()
    assert got["counts"][535:540].tolist() == want[:5] and got["counts"][541:542].tolist() == want[6:]
    assert int(got["counts"][540]) == 2 * ((1 << 60) + 12345) + 1
    assert abs(float(got["sums"][0]) - float(o["rel1"][ok].sum()) * 1e-6) < 1e-6


def test_wire_format_codecs():
    """felt252 <-> wsad codecs and calldata layouts (client/contract.py:35-53,
    contract/src/math.cairo:22-50, tests/test_math.cairo:95-96, contract/README.md:40-66)."""
    import numpy as np
    from svoc import client as c
    P = c.UPPER_BOUND_FELT252
    assert c.felt_to_i128(P - 1) == -1 and c.i128_to_felt(-1) == P - 1          # test_math.cairo:95-96
    x = np.array([[0.0, 0.4999999, 0.5, 1.0], [-0.25, -1e-7, 0.123456789, 20.714]])
    w = c.float_to_wsad_array(x)
    assert w.tolist() == [[int(v * 1e6) for v in row] for row in x.tolist()]      # truncation toward zero
    felts = c.wsad_array_to_felts(w)
    assert felts[1][0] == P - 250000 and felts[0][3] == 1000000
    assert [[c.float_to_fwsad(v) for v in row] for row in x.tolist()] == felts      # the scalar client helper
    back = c.felts_to_wsad_array(felts)
    assert back.dtype == np.int64 and np.array_equal(back, w)
    assert np.array_equal(c.felts_to_wsad_array([[c.to_hex(f) for f in row] for row in felts]), w)
    np.testing.assert_allclose(c.wsad_to_float_array(back), np.trunc(x * 1e6) * 1e-6)
    assert [c.fwsad_to_float(f) for f in felts[1]] == c.wsad_to_float_array(back)[1].tolist()
    import torch
    assert torch.equal(c.float_to_wsad_array(torch.from_numpy(x)), torch.from_numpy(w))
    assert c.update_prediction_calldata([0.5, 0.25]) == [2, 500000, 250000]
    cd = c.constructor_calldata(["0x1", "0x2", "0x3"], True, 2, 2, True, 0, 6, list(range(10, 17)))
    assert cd == [3, 1, 2, 3, 1, 2, 2, 1, 0, 6, 7, 10, 11, 12, 13, 14, 15, 16]        # contract/README.md:40-66
    assert c.constructor_calldata([5], False, 1, 0, False, -3, 1, [9])[6] == P - 3


def test_32_bit_square_reduction_is_exact_over_the_constrained_range():
    """csrc/svoc_common.cuh wmul_sq_small: wsad_mul(d, d) = (d*d + 500000) / 1000000 (signed_decimal.cairo:110-112,
    math.cairo:170-173) for |d| <= 1 048 999 with 32-bit intermediates only — the restated formula, checked here
    against plain 64-bit arithmetic for EVERY value of the range, and the 32-bit bound on its numerator."""
    a = np.arange(0, 1_049_000, dtype=np.int64)
    a1, a0 = a // 1000, a % 1000
    num = 2000 * a1 * a0 + a0 * a0 + 500000
    assert int(num.max()) < 2**31
    assert np.array_equal(a1 * a1 + num // 1_000_000, (a * a + 500000) // 1_000_000)
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "stochastic-vector-oracle-consensus_b200", "csrc", "svoc_common.cuh")).read()
    assert "a1 * a1 + (2000u * a1 * a0 + a0 * a0 + 500000u) / 1000000u" in src      # the formula tested is the one compiled


def test_multiply_high_division_is_exact_inside_its_stated_range():
    """csrc/generate_device.cuh UniformDiv: q = umulhi(e, floor(2^32 / d) + 1) equals e // d whenever e * d < 2^32
    (the `exact` flag; outside it the kernel divides).  Checked on every e for small divisors and on the edges of the
    range for large ones."""
    rng = np.random.default_rng(5)
    divisors = list(range(2, 70)) + [127, 128, 255, 256, 257, 1000, 4096, 14, 32768, 65535, 65536] + rng.integers(2, 1 << 16, 200).tolist()
    for d in divisors:
        m = (0xFFFFFFFF // d + 1) & 0xFFFFFFFF
        top = (1 << 32) // d                                  # e * d < 2^32  <=>  e < 2^32 / d
        if top <= 1 << 20:
            e = np.arange(0, top, dtype=np.uint64)
        else:
            e = np.unique(np.concatenate([np.arange(0, 1 << 16, dtype=np.uint64), rng.integers(0, top, 1 << 16).astype(np.uint64),
                                          np.arange(max(top - (1 << 12), 0), top, dtype=np.uint64),
                                          (np.arange(1, min(top // d, 1 << 12), dtype=np.uint64) * d), (np.arange(1, min(top // d, 1 << 12), dtype=np.uint64) * d - 1)]))
            e = e[e < top]
        assert np.array_equal((e * np.uint64(m)) >> np.uint64(32), e // np.uint64(d)), d
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "stochastic-vector-oracle-consensus_b200", "csrc", "generate_device.cuh")).read()
    assert "m(dd > 1u ? 0xffffffffu / dd + 1u : 0u)" in src and "(u64)max_e * dd < 0x100000000ull" in src


def test_philox_round_keys_are_the_key_schedule():
    """csrc/philox.cuh philox_round_keys: round r uses seed + r * (0x9E3779B9, 0xBB67AE85) — the schedule the pinned
    Python Philox (tests/philox_ref.py, checked against the Random123 known answers) applies inside its loop."""
    from philox_ref import philox4x32_10
    seed = 0x0123_4567_89AB_CDEF
    k0, k1 = seed & 0xFFFFFFFF, seed >> 32
    keys = [((k0 + r * 0x9E3779B9) & 0xFFFFFFFF, (k1 + r * 0xBB67AE85) & 0xFFFFFFFF) for r in range(10)]

    def with_round_keys(c):
        c0, c1, c2, c3 = c
        for a, b in keys:
            p0, p1 = 0xD2511F53 * c0, 0xCD9E8D57 * c2
            c0, c1, c2, c3 = (p1 >> 32) ^ c1 ^ a, p1 & 0xFFFFFFFF, (p0 >> 32) ^ c3 ^ b, p0 & 0xFFFFFFFF
        return (c0, c1, c2, c3)

    for c in [(0, 0, 0, 0), (1, 2, 3, 4), (0xFFFFFFFF, 0x12345678, 0x9ABCDEF0, 0x20000003)]:
        assert with_round_keys(c) == tuple(philox4x32_10(c, (k0, k1)))
