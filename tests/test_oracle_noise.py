"""Oracle: Philox known answers, table / layout / offset spec (CPU)."""
import numpy as np

from oracle import noise


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32 10 rounds (Salmon et al., SC'11)
    kats = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
            ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
            ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
             (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kats:
        got = tuple(int(x) for x in noise.philox4x32_10(*ctr, *key))
        assert got == want


def test_table_is_standard_normal_and_sliceable():
    t = noise.table_slice(1234, 0, 1 << 18)
    assert t.dtype == np.float16
    f = t.astype(np.float64)
    assert abs(f.mean()) < 0.01 and abs(f.std() - 1.0) < 0.01
    assert np.array_equal(noise.table_slice(1234, 1001, 777), t[1001:1778])
    assert not np.array_equal(noise.table_slice(1235, 0, 64), t[:64])


def test_layout_matches_reference_counts():
    lay = noise.flat_layout(9487, 512, 512, 2048)
    assert lay[-1][2] + lay[-1][3] == 13400848
    assert dict((n, s) for n, _, s, _ in lay)['core.i2h.weight'] == 10774288
    lay = noise.flat_layout(9487, 128, 128, 2048)
    assert lay[-1][2] + lay[-1][3] == 2865808      # src/algorithm/tools/utils.py:180
    lay = noise.flat_layout(9487, 128, 128, 2048, True, True)
    assert lay[-1][2] + lay[-1][3] == 2868624
    G = noise.granularity(512, 512, 2048)
    play, D, D_pad = noise.padded_layout(noise.flat_layout(9487, 512, 512, 2048), G)
    assert G == 2048 and D == 13400848 and D_pad % G == 0
    for name, shape, start, cnt, pstart in play:
        assert pstart % G == 0


def test_offsets_are_slot_aligned_deterministic_and_in_range(golden, golden_mod):
    for name, c in golden_mod.CASES.items():
        dims, G, play, D, D_pad, L, theta, fc, offs = golden_mod.case_inputs(c)
        assert np.array_equal(offs, golden[name + '_offsets'])
        assert (offs % G == 0).all() and (offs >= 0).all() and (offs + D_pad <= L).all()
    a = noise.sample_offsets(1, 0, 1000, 1 << 24, 75904, 128)
    b = noise.sample_offsets(1, 1, 1000, 1 << 24, 75904, 128)
    assert not np.array_equal(a, b) and len(np.unique(a)) > 900


def test_empty_and_single():
    assert noise.sample_offsets(1, 0, 0, 1 << 20, 1024, 64).shape == (0,)
    assert noise.table_slice(1, 5, 0).shape == (0,)
