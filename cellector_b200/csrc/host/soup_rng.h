// soup_rng.h — the random draws of souporcell's initialisation (souporcell/souporcell/src/main.rs:62-67, 354-364, 623-629).
// The reference uses rand's `StdRng` — rand 0.7.2 with rand_chacha 0.2.1 and rand_core 0.5.1, pinned at
// souporcell/souporcell/Cargo.lock:319-364 (Cargo.toml: rand = "0.7.0") — third-party crates that are NOT in
// /root/reference.  Their published definition is restated here:
//   StdRng (rand 0.7)            = ChaCha with 20 rounds (rand_chacha 0.2 ChaCha20Rng), seeded with 32 key bytes,
//                                  64-bit block counter from 0 (state words 12-13), stream id 0 (words 14-15);
//                                  the output is the sequence of 32-bit little-endian words of consecutive blocks;
//   rng.gen::<u8>()              = next_u32() as u8;
//   rng.gen::<f32>()             = (next_u32() >> 8) as f32 * 2^-24     (Standard: 24 random bits in [0, 1)).
// The ChaCha block function is checked against the RFC 7539 §2.3.2 vector (souporcell --rng_selftest,
// tests/test_cli_host.py); the mapping from words to u8 / f32 and the seeding order cannot be checked here (no Rust
// toolchain, no vectors in the repository): PARITY UNPINNED for the draws.  The EM that follows does not depend on it.
#pragma once
#include <cstdint>
#include <cstring>

namespace clxhost {

inline uint32_t rotl32(uint32_t x, int n) { return (x << n) | (x >> (32 - n)); }

// one ChaCha20 block: out = permuted(in) + in
inline void chacha20_block(const uint32_t in[16], uint32_t out[16]) {
    uint32_t x[16];
    memcpy(x, in, sizeof(x));
#define CLX_QR(a, b, c, d)                 \
    x[a] += x[b]; x[d] = rotl32(x[d] ^ x[a], 16); \
    x[c] += x[d]; x[b] = rotl32(x[b] ^ x[c], 12); \
    x[a] += x[b]; x[d] = rotl32(x[d] ^ x[a], 8);  \
    x[c] += x[d]; x[b] = rotl32(x[b] ^ x[c], 7);
    for (int round = 0; round < 10; ++round) {  // 10 double rounds
        CLX_QR(0, 4, 8, 12) CLX_QR(1, 5, 9, 13) CLX_QR(2, 6, 10, 14) CLX_QR(3, 7, 11, 15)
        CLX_QR(0, 5, 10, 15) CLX_QR(1, 6, 11, 12) CLX_QR(2, 7, 8, 13) CLX_QR(3, 4, 9, 14)
    }
#undef CLX_QR
    for (int i = 0; i < 16; ++i) out[i] = x[i] + in[i];
}

struct StdRng07 {
    uint32_t state[16];
    uint32_t buf[16];
    int pos = 16;

    static StdRng07 from_seed(const uint8_t seed[32]) {
        StdRng07 r;
        r.state[0] = 0x61707865u;  // "expand 32-byte k"
        r.state[1] = 0x3320646eu;
        r.state[2] = 0x79622d32u;
        r.state[3] = 0x6b206574u;
        for (int i = 0; i < 8; ++i)
            r.state[4 + i] = (uint32_t)seed[4 * i] | (uint32_t)seed[4 * i + 1] << 8 | (uint32_t)seed[4 * i + 2] << 16 | (uint32_t)seed[4 * i + 3] << 24;
        r.state[12] = r.state[13] = 0;  // 64-bit block counter
        r.state[14] = r.state[15] = 0;  // stream id
        r.pos = 16;
        return r;
    }
    uint32_t next_u32() {
        if (pos == 16) {
            chacha20_block(state, buf);
            if (++state[12] == 0) ++state[13];
            pos = 0;
        }
        return buf[pos++];
    }
    uint8_t gen_u8() { return (uint8_t)next_u32(); }
    float gen_f32() { return (float)(next_u32() >> 8) * (1.0f / 16777216.0f); }
    // new_seed (main.rs:623-629)
    void new_seed(uint8_t out[32]) {
        for (int i = 0; i < 32; ++i) out[i] = gen_u8();
    }
};

}  // namespace clxhost
