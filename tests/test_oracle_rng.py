"""oracle/rng.c against cuRAND's own XORWOW (golden from the toolkit's host tables,
tests/golden/make_xorwow_golden.cu): curand_init(seed, subsequence, offset) state and
the first four curand_uniform draws, bit-exact."""
import ctypes as C
import os

import numpy as np

import oraclebind as ob

GOLD = os.path.join(os.path.dirname(__file__), "golden", "xorwow_golden.txt")


def test_xorwow_state_and_uniforms_bit_exact():
    lib = ob.load()
    n = 0
    for line in open(GOLD):
        f = line.split()
        seed, sub, off = int(f[0]), int(f[1]), int(f[2])
        st = ob.rng_state(seed, sub, off)
        assert list(st) == [int(x) for x in f[3:9]], (seed, sub, off)
        for k in range(4):
            u = np.float32(lib.orc_rng_uniform(C.byref(st)))
            assert u.view(np.uint32) == int(f[9 + k], 16), (seed, sub, off, k)
        n += 1
    assert n >= 200


def test_image_init_matches_per_pixel_seeding():
    img = ob.rng_init_image(1234, 40, 7)
    for (y, x) in [(0, 0), (3, 17), (6, 39)]:
        assert list(img[y, x]) == list(ob.rng_state(1234, y, x))
