"""Pure-Python restatement of ``_generate_positions`` (TEST INFRASTRUCTURE ONLY).

Follows reference src/simulation.rs:107-166 and the published algorithms of the
crates it calls (rand 0.9.2 / rand_chacha 0.9.0 / rand_core 0.9.5, versions
pinned in reference Cargo.lock:1739-1810):

* ``ChaCha8Rng::seed_from_u64``: the 32-byte key is filled by a PCG32 stream
  (rand_core ``SeedableRng::seed_from_u64``);
* ChaCha with 8 rounds, 64-bit block counter starting at 0, stream 0; the block
  words are consumed in order (``next_u32``);
* ``random_range`` for f32: 23 random mantissa bits -> [1,2) - 1 -> ``v*scale+low``
  (``UniformFloat::sample_single``), retrying with a smaller scale if the result
  reaches ``high``;
* ``random_range(..n)`` for usize < 2^32: Canon's widening-multiply method on u32
  (``UniformInt<u32>::sample_single_inclusive``);
* ``IteratorRandom::choose_multiple``: reservoir sampling.

PINNED by the reference's own golden vectors (src/simulation.rs:168-219), see
tests/test_generate_positions.py.
"""

from __future__ import annotations

import numpy as np

import ctypes
import ctypes.util

# Rust's f32::cos / f32::sin call the platform libm (glibc cosf/sinf on the reference's Linux targets);
# numpy's SIMD float32 kernels differ from it by an ulp on some arguments.
_libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
for _f in (_libm.cosf, _libm.sinf):
    _f.restype = ctypes.c_float
    _f.argtypes = [ctypes.c_float]


def _cosf(x):
    return np.float32(_libm.cosf(float(x)))


def _sinf(x):
    return np.float32(_libm.sinf(float(x)))


M32 = 0xFFFFFFFF
M64 = 0xFFFFFFFFFFFFFFFF
f32 = np.float32


def _rotl(x, n):
    return ((x << n) & M32) | (x >> (32 - n))


class ChaCha8:
    def __init__(self, seed_u64: int):
        state = seed_u64 & M64
        key = []
        for _ in range(8):
            state = (state * 6364136223846793005 + 11634580027462260723) & M64
            xorshifted = (((state >> 18) ^ state) >> 27) & M32
            rot = state >> 59
            key.append(((xorshifted >> rot) | (xorshifted << ((32 - rot) & 31))) & M32 if rot else xorshifted)
        self.key = key
        self.counter = 0
        self.buf = []

    def _block(self):
        c = self.counter
        init = [0x61707865, 0x3320646E, 0x79622D32, 0x6B206574] + self.key + [c & M32, (c >> 32) & M32, 0, 0]
        x = list(init)

        def qr(a, b, cc, d):
            x[a] = (x[a] + x[b]) & M32; x[d] = _rotl(x[d] ^ x[a], 16)
            x[cc] = (x[cc] + x[d]) & M32; x[b] = _rotl(x[b] ^ x[cc], 12)
            x[a] = (x[a] + x[b]) & M32; x[d] = _rotl(x[d] ^ x[a], 8)
            x[cc] = (x[cc] + x[d]) & M32; x[b] = _rotl(x[b] ^ x[cc], 7)

        for _ in range(4):  # 8 rounds = 4 double rounds
            qr(0, 4, 8, 12); qr(1, 5, 9, 13); qr(2, 6, 10, 14); qr(3, 7, 11, 15)
            qr(0, 5, 10, 15); qr(1, 6, 11, 12); qr(2, 7, 8, 13); qr(3, 4, 9, 14)
        self.counter += 1
        return [(a + b) & M32 for a, b in zip(x, init)]

    def next_u32(self) -> int:
        if not self.buf:
            self.buf = self._block()
        return self.buf.pop(0)

    def range_f32(self, low, high):
        low, high = f32(low), f32(high)
        assert low < high
        scale = f32(high - low)
        while True:
            bits = (self.next_u32() >> 9) | 0x3F800000
            v12 = np.array([bits], dtype=np.uint32).view(np.float32)[0]
            v01 = f32(v12 - f32(1.0))
            res = f32(f32(v01 * scale) + low)
            if res < high:
                return res
            # decrease scale by one ulp and retry
            scale = np.array([np.array([scale], np.float32).view(np.uint32)[0] - 1], np.uint32).view(np.float32)[0]

    def below_usize(self, n: int) -> int:
        """random_range(..n) for n <= 2^32."""
        rng = n & M32
        if rng == 0:
            return self.next_u32()
        m = self.next_u32() * rng
        result, lo = m >> 32, m & M32
        if lo > ((-rng) & M32):
            new_hi = (self.next_u32() * rng) >> 32
            if lo + new_hi > M32:
                result += 1
        return result


def generate_positions(n_agents, spring_length, domain_size, domain_height, rng_seed=0, dx=(0.0, 0.0),
                       randomize_positions=0.0, n_vertices=8):
    rng = ChaCha8(rng_seed)
    spring_length = f32(spring_length)
    s = f32(min(max(f32(randomize_positions), f32(0.0)), f32(1.0)))
    n_chunk_sides = int(np.ceil(np.sqrt(f32(n_agents))))
    dx0, dx1 = f32(dx[0]), f32(dx[1])
    dchunk1 = f32(f32(f32(domain_size[0]) - f32(f32(2.0) * dx0)) / f32(n_chunk_sides))
    dchunk2 = f32(f32(f32(domain_size[1]) - f32(f32(2.0) * dx1)) / f32(n_chunk_sides))
    all_indices = [(i, j) for i in range(n_chunk_sides) for j in range(n_chunk_sides)]
    # IteratorRandom::choose_multiple
    picked = all_indices[:n_agents]
    if len(picked) == n_agents:
        for i, elem in enumerate(all_indices[n_agents:]):
            k = rng.below_usize(i + 1 + n_agents)
            if k < len(picked):
                picked[k] = elem
    half = f32(f32(f32(n_vertices) * spring_length) / f32(2.0))
    H = f32(domain_height)

    def s_gen():
        if s == 0.0:
            return f32(1.0)
        return rng.range_f32(f32(1.0) - s, f32(1.0) + s)

    out = []
    for (i0, i1) in picked:
        xlow = f32(dx0 + f32(f32(i0) * dchunk1))
        ylow = f32(dx1 + f32(f32(i1) * dchunk2))
        mid = [
            rng.range_f32(xlow + half, f32(xlow + dchunk1) - half),
            rng.range_f32(ylow + half, f32(ylow + dchunk2) - half),
            rng.range_f32(f32(0.4) * H, f32(0.6) * H),
        ]
        angle = rng.range_f32(f32(0.0), f32(f32(2.0) * f32(np.pi)))
        p1 = [f32(mid[0] - f32(half * _cosf(angle))),
              f32(mid[1] - f32(half * _sinf(angle))),
              f32(mid[2] - f32(half * f32(0.0)))]
        pos = np.zeros((n_vertices, 3), dtype=np.float32)
        for c in range(3):  # nalgebra from_fn is column-major
            for r in range(n_vertices):
                a = f32(f32(f32(r) * spring_length) * s_gen())
                if c == 0:
                    b = _cosf(f32(angle * s_gen()))
                elif c == 1:
                    b = _sinf(f32(angle * s_gen()))
                else:
                    b = f32(0.0)
                pos[r, c] = f32(p1[c] + f32(a * b))
        out.append(pos)
    return out
