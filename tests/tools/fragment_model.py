"""Lane-level numpy model of the decode kernel's register/smem index maps (csrc/decode.cu).

Not a test of the CUDA code itself: it replays, lane by lane, the address arithmetic the consumer warps use
(128-byte TMA swizzle, mma.sync m16n8k16 fragment ownership, the PRMT byte gather for fp8 V, LDSM.T for
16-bit V) and checks that the assembled matrices equal Q.K^T and P.V.  Run on CPU before spending GPU time.
"""
import numpy as np

rng = np.random.default_rng(0)
D = 128


def swizzle_store(chunk_rows):  # [16][128 bytes] -> physical 2 KiB
    phys = np.zeros(16 * 128, dtype=chunk_rows.dtype)
    for r in range(16):
        for c in range(8):
            phys[r * 128 + ((c ^ (r & 7)) << 4): r * 128 + ((c ^ (r & 7)) << 4) + 16] = chunk_rows[r, 16 * c:16 * c + 16]
    return phys


def mma(A_frag, B_frag):
    """A_frag[lane] = 4 regs of 2 elems, B_frag[lane] = 2 regs of 2 elems -> C[16][8] from the PTX layout."""
    A = np.zeros((16, 16)); B = np.zeros((16, 8))
    for lane in range(32):
        g, t = lane >> 2, lane & 3
        a, b = A_frag[lane], B_frag[lane]
        A[g, 2 * t], A[g, 2 * t + 1] = a[0]
        A[g + 8, 2 * t], A[g + 8, 2 * t + 1] = a[1]
        A[g, 2 * t + 8], A[g, 2 * t + 9] = a[2]
        A[g + 8, 2 * t + 8], A[g + 8, 2 * t + 9] = a[3]
        B[2 * t, g], B[2 * t + 1, g] = b[0]
        B[2 * t + 8, g], B[2 * t + 9, g] = b[1]
    return A @ B


def c_frag(C, lane):
    g, t = lane >> 2, lane & 3
    return [C[g, 2 * t], C[g, 2 * t + 1], C[g + 8, 2 * t], C[g + 8, 2 * t + 1]]


def check_fp8():
    Q = rng.integers(-4, 5, size=(16, D)).astype(float)
    K = rng.integers(-4, 5, size=(16, D)).astype(float)   # "bytes" modelled as small ints
    V = rng.integers(-4, 5, size=(16, D)).astype(float)
    kph, vph = swizzle_store(K), swizzle_store(V)
    # ---- S = Q K^T ----
    SA = np.zeros((16, 8)); SB = np.zeros((16, 8))
    for st in range(8):
        Af, BfA, BfB = [], [], []
        for lane in range(32):
            g, t = lane >> 2, lane & 3
            dbase = 32 * t + 4 * st
            Af.append([(Q[g, dbase], Q[g, dbase + 1]), (Q[g + 8, dbase], Q[g + 8, dbase + 1]),
                       (Q[g, dbase + 2], Q[g, dbase + 3]), (Q[g + 8, dbase + 2], Q[g + 8, dbase + 3])])
            i, r = st >> 2, st & 3
            chunk = 2 * t + i
            for rows, dst in ((g, BfA), (g + 8, BfB)):
                addr = rows * 128 + ((chunk ^ g) << 4) + 4 * r
                wd = kph[addr:addr + 4]
                dst.append([(wd[0], wd[1]), (wd[2], wd[3])])
        SA += mma(Af, BfA); SB += mma(Af, BfB)
    S = np.concatenate([SA, SB], axis=1)      # [16 heads][16 tokens]
    assert np.array_equal(S, Q @ K.T), "fp8 QK map wrong"
    # bank check of the K loads: per quarter-warp the 8 lanes must hit 8 distinct 16B columns
    for i in range(2):
        for q in range(4):
            cols = set()
            for lane in range(8 * q, 8 * q + 8):
                g, t = lane >> 2, lane & 3
                cols.add(((2 * t + i) ^ g))
            assert len(cols) == 8, "K bank conflict"
    # ---- O^T = V^T P^T ----
    P = rng.integers(0, 4, size=(16, 16)).astype(float)  # [head][tok]
    OT = {}
    for r in range(4):
        for hb in range(2):
            Af, Bf0, Bf1 = [], [], []
            for lane in range(32):
                g, t = lane >> 2, lane & 3
                def vword(j):
                    addr = j * 128 + ((g ^ (j & 7)) << 4) + 4 * r
                    return vph[addr:addr + 4]
                w0, w1, w2, w3 = vword(2 * t), vword(2 * t + 1), vword(2 * t + 8), vword(2 * t + 9)
                def prmt(a, b, sel):
                    src = list(a) + list(b)
                    return [src[(sel >> (4 * k)) & 0xf] for k in range(4)]
                sel = 0x7362 if hb else 0x5140
                x01, x23 = prmt(w0, w1, sel), prmt(w2, w3, sel)
                Af.append([(x01[0], x01[1]), (x01[2], x01[3]), (x23[0], x23[1]), (x23[2], x23[3])])
                Bf0.append([(P[g, 2 * t], P[g, 2 * t + 1]), (P[g, 2 * t + 8], P[g, 2 * t + 9])])
                Bf1.append([(P[g + 8, 2 * t], P[g + 8, 2 * t + 1]), (P[g + 8, 2 * t + 8], P[g + 8, 2 * t + 9])])
            OT[(0, 2 * r + hb)] = mma(Af, Bf0); OT[(1, 2 * r + hb)] = mma(Af, Bf1)
    O = np.zeros((16, D))
    for (a, mt), C in OT.items():
        for lane in range(32):
            g, t = lane >> 2, lane & 3
            c = c_frag(C, lane)
            for hh in range(2):
                head = a * 8 + 2 * t + hh
                dlo = 16 * g + 2 * mt
                O[head, dlo] = c[hh]; O[head, dlo + 1] = c[2 + hh]
    assert np.array_equal(O, P @ V), "fp8 PV map wrong"
    for q in range(4):  # V bank check
        for rowsel in range(4):
            cols = set()
            for lane in range(8 * q, 8 * q + 8):
                g, t = lane >> 2, lane & 3
                j = [2 * t, 2 * t + 1, 2 * t + 8, 2 * t + 9][rowsel]
                cols.add(g ^ (j & 7))
            assert len(cols) == 8, "V bank conflict"


def check_16bit():
    Q = rng.integers(-4, 5, size=(16, D)).astype(float)
    K = rng.integers(-4, 5, size=(16, D)).astype(float)
    V = rng.integers(-4, 5, size=(16, D)).astype(float)
    # 16-bit chunk: two boxes [hf][16 rows][64 elems]; model one "byte" = half an element -> use element units:
    # physical index in elements: hf*1024 + row*64 + ((c ^ (row&7))*8) + e, c = 16B chunk (8 elems)
    def store(X):
        ph = np.zeros(2 * 16 * 64)
        for hf in range(2):
            for r in range(16):
                for c in range(8):
                    ph[hf * 1024 + r * 64 + ((c ^ (r & 7)) * 8): hf * 1024 + r * 64 + ((c ^ (r & 7)) * 8) + 8] = X[r, 64 * hf + 8 * c: 64 * hf + 8 * c + 8]
        return ph
    kph, vph = store(K), store(V)
    SA = np.zeros((16, 8)); SB = np.zeros((16, 8))
    for hf in range(2):
        for i in range(2):
            for sub in range(2):
                st = hf * 4 + i * 2 + sub
                Af, BfA, BfB = [], [], []
                for lane in range(32):
                    g, t = lane >> 2, lane & 3
                    dbase = 64 * (st >> 2) + 16 * t + 4 * (st & 3)
                    Af.append([(Q[g, dbase], Q[g, dbase + 1]), (Q[g + 8, dbase], Q[g + 8, dbase + 1]),
                               (Q[g, dbase + 2], Q[g, dbase + 3]), (Q[g + 8, dbase + 2], Q[g + 8, dbase + 3])])
                    co = ((2 * t + i) ^ g) * 8
                    for rows, dst in ((g, BfA), (g + 8, BfB)):
                        base = hf * 1024 + rows * 64 + co
                        k8 = kph[base: base + 8]
                        w = k8[4 * sub: 4 * sub + 4]   # .x,.y for sub 0 ; .z,.w for sub 1
                        dst.append([(w[0], w[1]), (w[2], w[3])])
                SA += mma(Af, BfA); SB += mma(Af, BfB)
    S = np.concatenate([SA, SB], axis=1)
    assert np.array_equal(S, Q @ K.T), "16-bit QK map wrong"
    P = rng.integers(0, 4, size=(16, 16)).astype(float)
    O = np.zeros((16, D))
    for mt in range(8):
        # ldmatrix.x4.trans: lane L supplies the row address of matrix L>>3, row L&7
        mats = []
        for mi in range(4):
            M = np.zeros((8, 8))
            for rr in range(8):
                tokrow = (mi >> 1) * 8 + rr
                chunk = 2 * (mt & 3) + (mi & 1)
                base = (mt >> 2) * 1024 + tokrow * 64 + ((chunk ^ rr) * 8)
                M[rr] = vph[base: base + 8]
            mats.append(M)
        Af, Bf0, Bf1 = [], [], []
        for lane in range(32):
            g, t = lane >> 2, lane & 3
            # .trans: thread (g,t) receives M[2t][g], M[2t+1][g]
            Af.append([(M[2 * t, g], M[2 * t + 1, g]) for M in mats])
            Bf0.append([(P[g, 2 * t], P[g, 2 * t + 1]), (P[g, 2 * t + 8], P[g, 2 * t + 9])])
            Bf1.append([(P[g + 8, 2 * t], P[g + 8, 2 * t + 1]), (P[g + 8, 2 * t + 8], P[g + 8, 2 * t + 9])])
        for a, Bf in ((0, Bf0), (1, Bf1)):
            C = mma(Af, Bf)
            for lane in range(32):
                g, t = lane >> 2, lane & 3
                c = c_frag(C, lane)
                for hh in range(2):
                    head = a * 8 + 2 * t + hh
                    O[head, 16 * mt + g] = c[hh]; O[head, 16 * mt + g + 8] = c[2 + hh]
    assert np.array_equal(O, P @ V), "16-bit PV map wrong"


def check_swapped_qk_and_movmatrix():
    """Round-1b layout: S^T(16 tok x 8 heads) = K.Q^T with K as the A operand, then movmatrix.trans to the P^T B operand."""
    Q = rng.integers(-4, 5, size=(8, D)).astype(float)       # 8 heads
    K = rng.integers(-4, 5, size=(16, D)).astype(float)
    kph = swizzle_store(K)
    ST = np.zeros((16, 8))
    for st in range(8):
        Af, Bf = [], []
        for lane in range(32):
            g, t = lane >> 2, lane & 3
            dbase = 32 * t + 4 * st
            i, r = st >> 2, st & 3
            chunk = 2 * t + i
            wg = kph[g * 128 + ((chunk ^ g) << 4) + 4 * r:][:4]
            wg8 = kph[(g + 8) * 128 + ((chunk ^ g) << 4) + 4 * r:][:4]
            # a0 = lo(word row g), a1 = lo(word row g+8), a2 = hi(row g), a3 = hi(row g+8)
            Af.append([(wg[0], wg[1]), (wg8[0], wg8[1]), (wg[2], wg[3]), (wg8[2], wg8[3])])
            Bf.append([(Q[g, dbase], Q[g, dbase + 1]), (Q[g, dbase + 2], Q[g, dbase + 3])])
        ST += mma(Af, Bf)
    assert np.array_equal(ST, K @ Q.T), "swapped QK map wrong"
    # C fragment -> packed pairs -> movmatrix.trans -> B fragment of P^T (k = token, n = head)
    P = ST  # treat logits as probabilities for the layout check: P^T[tok][head]
    r0 = {}; r1 = {}
    for lane in range(32):
        g, t = lane >> 2, lane & 3
        c = c_frag(P, lane)          # c0,c1 = [tok g][head 2t,2t+1]; c2,c3 = [tok g+8][...]
        r0[lane] = (c[0], c[1]); r1[lane] = (c[2], c[3])
    def movm(regs):
        M = np.zeros((8, 8))
        for lane in range(32):
            g, t = lane >> 2, lane & 3
            M[g, 2 * t], M[g, 2 * t + 1] = regs[lane]
        return {lane: (M[2 * (lane & 3), lane >> 2], M[2 * (lane & 3) + 1, lane >> 2]) for lane in range(32)}
    b0, b1 = movm(r0), movm(r1)
    for lane in range(32):
        g, t = lane >> 2, lane & 3
        assert b0[lane] == (P[2 * t, g], P[2 * t + 1, g])          # {B[k=2t][n=g], B[k=2t+1][n=g]}, B[k][n] = P^T[tok k][head n]
        assert b1[lane] == (P[2 * t + 8, g], P[2 * t + 9, g])


if __name__ == "__main__":
    check_swapped_qk_and_movmatrix()
    check_fp8(); check_16bit(); print("fragment maps OK")
