"""The identity k_seed_multi uses to get the reverse strand's packed words from the forward strand's
(fem_b200/csrc/k_seed_multi.cuh, "both strands of a read at once"): a 16-base window of the forward 2-bit
stream, its 2-bit groups reversed (bit reversal + swap of neighbouring bits), complemented, is the packed
word of the reverse complement (src/sequence_batch.h:90-98 builds that strand base by base).  Checked here
in plain Python against packing the reverse complement directly; bases past the read's end are "don't care"
(no q-gram reads them), ambiguous bases are hashed as A on both strands (src/utils.h:95-100)."""
import random


def pack(codes, n_words):
    words = [0] * n_words
    for i, c in enumerate(codes):
        if c < 4:
            words[i >> 4] |= c << (30 - 2 * (i & 15))
    return words


def brev32(x):
    return int("{:032b}".format(x)[::-1], 2)


def funnelshift_l(lo, hi, sh):
    return (((hi << 32) | lo) << sh >> 32) & 0xFFFFFFFF


def reverse_words(fwd, l, n_words, codes):
    out = []
    for w in range(n_words):
        s0 = l - 16 - 16 * w
        a0, sh = s0 >> 4, 2 * (s0 & 15)  # Python's >> and & floor like the kernel's arithmetic shift / mask
        hi = fwd[a0] if 0 <= a0 < n_words else 0
        lo = fwd[a0 + 1] if 0 <= a0 + 1 < n_words else 0
        x = brev32(funnelshift_l(lo, hi, sh))
        x = ((x & 0x55555555) << 1) | ((x >> 1) & 0x55555555)
        out.append(~x & 0xFFFFFFFF)
    for i, c in enumerate(codes):  # ambiguous bases: cleared afterwards (atomicAnd in the kernel)
        if c == 4:
            j = l - 1 - i
            out[j >> 4] &= ~(3 << (30 - 2 * (j & 15))) & 0xFFFFFFFF
    return out


def test_reverse_strand_words_from_forward_words():
    rng = random.Random(20260)
    for l in list(range(1, 70)) + [100, 127, 128, 129, 150, 250, 255, 256, 257, 1000, 1024]:
        for amb in (0.0, 0.05):
            codes = [4 if rng.random() < amb else rng.randrange(4) for _ in range(l)]
            n_words = (l + 15) // 16 + 1
            fwd = pack(codes, n_words)
            rc = [4 if c == 4 else 3 - c for c in reversed(codes)]
            want = pack(rc, n_words)
            got = reverse_words(fwd, l, n_words, codes)
            for j in range(l):  # every base a q-gram can cover
                w, sh = j >> 4, 30 - 2 * (j & 15)
                assert (got[w] >> sh) & 3 == (want[w] >> sh) & 3, (l, amb, j)
