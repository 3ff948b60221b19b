"""Seeded mutations of JPEG files for the robustness tests (same families as tests/hostsim/fuzz_main.cpp)."""
import random


def scan_start(f):
    i = 2
    while i + 3 < len(f):
        if f[i] != 0xFF:
            return 0
        l = (f[i + 2] << 8) | f[i + 3]
        if f[i + 1] == 0xDA:
            return i + 2 + l
        i += 2 + l
    return 0


def mutate(data, rng, in_scan):
    f = bytearray(data)
    for _ in range(1 + rng.randrange(3)):
        ss = scan_start(f)
        lo, hi = 2, len(f)
        if ss and ss < len(f):
            if in_scan:
                lo = ss
            else:
                hi = ss
        if hi <= lo:
            break
        at = lo + rng.randrange(hi - lo)
        kind = rng.randrange(7)
        if kind == 0:
            f[at] ^= 1 << rng.randrange(8)
        elif kind == 1:
            f[at] = rng.randrange(256)
        elif kind == 2:
            f[at] = 0xFF
        elif kind == 3:
            del f[at : at + 1 + rng.randrange(400)]
        elif kind == 4:
            f[at:at] = bytes(rng.randrange(256) for _ in range(1 + rng.randrange(64)))
        elif kind == 5:
            del f[at:]
        else:
            v = 0x00 if rng.randrange(2) else 0xFF
            for i in range(at, min(len(f), at + 1 + rng.randrange(32))):
                f[i] = v
    return bytes(f)


def cases(seeds, n, seed, scan_share=0.75):
    """n mutated files drawn from the `seeds` list of bytes objects."""
    rng = random.Random(seed)
    return [mutate(seeds[rng.randrange(len(seeds))], rng, rng.random() < scan_share) for _ in range(n)]
