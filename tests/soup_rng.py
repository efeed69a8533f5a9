"""The draws of souporcell's initialisation in plain Python (test infrastructure): ChaCha20 as published (RFC 7539) and
rand 0.7's StdRng conventions as cellector_b200/csrc/host/soup_rng.h states them — an independent statement of the same
definition, so the C++ host and the tests derive the initial centers separately."""
import numpy as np

M = 0xFFFFFFFF


def _rotl(x, n):
    return ((x << n) & M) | (x >> (32 - n))


def chacha20_block(state):
    x = list(state)

    def qr(a, b, c, d):
        x[a] = (x[a] + x[b]) & M; x[d] = _rotl(x[d] ^ x[a], 16)
        x[c] = (x[c] + x[d]) & M; x[b] = _rotl(x[b] ^ x[c], 12)
        x[a] = (x[a] + x[b]) & M; x[d] = _rotl(x[d] ^ x[a], 8)
        x[c] = (x[c] + x[d]) & M; x[b] = _rotl(x[b] ^ x[c], 7)

    for _ in range(10):
        qr(0, 4, 8, 12); qr(1, 5, 9, 13); qr(2, 6, 10, 14); qr(3, 7, 11, 15)
        qr(0, 5, 10, 15); qr(1, 6, 11, 12); qr(2, 7, 8, 13); qr(3, 4, 9, 14)
    return [(a + b) & M for a, b in zip(x, state)]


class StdRng07:
    def __init__(self, seed: bytes):
        assert len(seed) == 32
        self.state = [0x61707865, 0x3320646E, 0x79622D32, 0x6B206574] + \
                     [int.from_bytes(seed[4 * i:4 * i + 4], "little") for i in range(8)] + [0, 0, 0, 0]
        self.buf, self.pos = [], 16

    def next_u32(self):
        if self.pos == 16:
            self.buf = chacha20_block(self.state)
            self.state[12] = (self.state[12] + 1) & M
            if self.state[12] == 0:
                self.state[13] = (self.state[13] + 1) & M
            self.pos = 0
        v = self.buf[self.pos]
        self.pos += 1
        return v

    def gen_u8(self):
        return self.next_u32() & 0xFF

    def gen_f32(self):
        return np.float32(self.next_u32() >> 8) * np.float32(1.0 / 16777216.0)

    def new_seed(self):
        return bytes(self.gen_u8() for _ in range(32))


def souporcell_initial_centers(seed: int, threads: int, restarts: int, k: int, n_loci: int):
    """[runs][k][n_loci] f32 in the order souporcell_main (main.rs:62-88) draws them, thread-major"""
    main = StdRng07(bytes([seed]) * 32)
    per_thread = -(-restarts // threads)
    out = np.empty((threads * per_thread, k, n_loci), np.float32)
    r = 0
    for _ in range(threads):
        rng = StdRng07(main.new_seed())
        for _ in range(per_thread):
            for q in range(k):
                for l in range(n_loci):
                    out[r, q, l] = min(max(rng.gen_f32(), np.float32(0.0001)), np.float32(0.9999))
            r += 1
    return out
