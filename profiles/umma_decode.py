"""Decodes which shared-memory word a tcgen05.mma (kind::tf32) treats as A[m][k] for a given matrix descriptor:
the A image holds its own word index, B is an identity on k, so D[m][k] = index of the word read.

    python profiles/umma_decode.py <variant>      one variant per process (a faulting descriptor kills the context)
    python profiles/umma_decode.py all            runs every variant in a subprocess and prints a summary
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

WORDS = 32768          # 128 KB image


def idesc(M, N, a_mn=0, b_mn=0):
    return (1 << 4) | (2 << 7) | (2 << 10) | (a_mn << 15) | (b_mn << 16) | ((N >> 3) << 17) | ((M >> 4) << 24)


def desc(start16=0, lbo=16, sbo=1024, layout=2, base_offset=0):
    return (start16 & 0x3FFF) | ((lbo >> 4) << 16) | ((sbo >> 4) << 32) | (1 << 46) | (base_offset << 49) | (layout << 61)


def b_identity(N):
    img = np.zeros((N, 8, 4), np.float32)
    for r in range(min(N, 8)):
        img[r, (r // 4) ^ (r & 7), r % 4] = 1.0          # K-major SWIZZLE_128B tile: B[n][k] = (n == k)
    return img.tobytes()


# name -> (a_mn, layout, lbo, sbo, start_bytes, base_offset, M)
VARIANTS = {
    "k_sw128_aligned": (0, 2, 16, 1024, 0, 0, 128),
    "k_sw128_row3": (0, 2, 16, 1024, 3 * 128, 0, 128),
    "k_sw128_row3_bo3": (0, 2, 16, 1024, 3 * 128, 3, 128),
    "k_none_row0": (0, 0, 4096, 128, 0, 0, 128),
    "k_none_row3": (0, 0, 4096, 128, 3 * 16, 0, 128),
    "mn_sw128_aligned": (1, 2, 8192, 1024, 0, 0, 64),
    "mn_sw128_row3": (1, 2, 8192, 1024, 3 * 128, 0, 64),
    "mn_sw128_row3_bo3": (1, 2, 8192, 1024, 3 * 128, 3, 64),
    "mn_sw128_aligned_m128": (1, 2, 8192, 1024, 0, 0, 128),
    "mn_none_a": (1, 0, 128, 4096, 0, 0, 64),
    "mn_none_b": (1, 0, 4096, 128, 0, 0, 64),
    "mn_sw32": (1, 6, 4096, 256, 0, 0, 64),
}


def one(name):
    from scala_akka_tensorflow_sharetrade_helper_b200 import _capi
    lib = _capi.lib()
    a_mn, layout, lbo, sbo, start, bo, M = VARIANTS[name]
    idx = np.arange(WORDS)
    res = []
    for part in (idx % 1024 + 1, idx // 1024 + 1):      # (+1: a zero result then means 'not read from the image')
        a_img = part.astype(np.float32).tobytes()
        out = np.empty((128, 16), np.float32)
        st = lib.qp_debug_umma(0, a_img, len(a_img), b_identity(16), 16 * 128, C.c_uint64(desc(start >> 4, lbo, sbo, layout, bo)),
                               C.c_uint64(desc()), idesc(M, 16, a_mn=a_mn), 1, 0, 0, 0, 0, 16, C.c_float(-1.0),
                               out.ctypes.data_as(C.c_void_p), 0, 0)
        if st != 0:
            print(name, "FAILED:", _capi.last_error())
            return
        res.append(out[:, :8])
    word = ((res[0] - 1) + 1024 * (res[1] - 1)).astype(np.int64)          # [lane][k] -> word index
    word[res[0] == 0] = -1                                                # product was zero: nothing read from the image
    print(f"== {name}: a_mn={a_mn} layout={layout} LBO={lbo} SBO={sbo} start={start} base_offset={bo} M={M}")
    lanes = [l for l in range(128) if res[0][l, 0] >= 0]
    print("   lanes written:", lanes[:20], "..." if len(lanes) > 20 else "", len(lanes))
    for l in lanes[:6] + lanes[8:10] + lanes[16:18] + lanes[32:34] + lanes[63:65]:
        print(f"   lane {l:3d}: byte addresses of k=0..7:", [int(w) * 4 if w >= 0 else None for w in word[l]], ' raw', res[0][l].tolist())


if __name__ == "__main__":
    if sys.argv[1] == "all":
        for v in VARIANTS:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), v], capture_output=True, text=True, timeout=120)
            print(r.stdout.strip() or f"== {v}: no output", flush=True)
            if r.returncode != 0:
                print("   rc", r.returncode, r.stderr.strip().splitlines()[-1:] if r.stderr else "")
    else:
        one(sys.argv[1])
