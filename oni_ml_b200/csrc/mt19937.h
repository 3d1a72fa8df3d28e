// mt19937.h -- the reference's random stream, host side.
//
// oni-lda-c seeds and draws through Cokus' 1998 recoding of MT19937
// (cokus.c:52-144): seed forced odd, state filled by x <- 69069*x mod 2^32,
// block regeneration of 624 words, standard tempering; myrand() divides a draw
// by 2^32 (lda-model.h:11).  "Same seed => same model" requires the same stream,
// so the generator is restated here in the textbook index form.
#pragma once
#include <stdint.h>

namespace oni {

class CokusMT {
  public:
    explicit CokusMT(uint32_t seed) { reseed(seed); }

    void reseed(uint32_t seed)
    {
        uint32_t x = seed | 1u;  // cokus.c:100
        for (int i = 0; i < kN; i++) {
            mt_[i] = x;
            x *= 69069u;  // cokus.c:103-104
        }
        idx_ = kN;  // first draw regenerates the block (left = 0, cokus.c:103)
    }

    uint32_t next()
    {
        if (idx_ >= kN) regenerate();
        uint32_t y = mt_[idx_++];
        y ^= y >> 11;
        y ^= (y << 7) & 0x9D2C5680u;
        y ^= (y << 15) & 0xEFC60000u;
        y ^= y >> 18;
        return y;
    }

    // lda-model.h:11
    double uniform() { return (double)next() / 4294967296.0; }

  private:
    static constexpr int kN = 624, kM = 397;
    uint32_t mt_[kN];
    int idx_;

    static uint32_t twist(uint32_t hi_src, uint32_t lo_src)
    {
        const uint32_t y = (hi_src & 0x80000000u) | (lo_src & 0x7FFFFFFFu);
        return (y >> 1) ^ ((lo_src & 1u) ? 0x9908B0DFu : 0u);
    }

    void regenerate()
    {
        for (int i = 0; i < kN; i++)
            mt_[i] = mt_[(i + kM) % kN] ^ twist(mt_[i], mt_[(i + 1) % kN]);
        idx_ = 0;
    }
};

}  // namespace oni
