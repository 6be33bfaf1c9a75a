#!/usr/bin/env python3
"""Executable specification of the BANDED Gotoh fill + traceback used by the GPU band kernels.

Claim (DESIGN.md "band DP"): let the band be the diagonals dlo <= c - r <= dhi (matrix coordinates, c = column =
position in A + 1, r = row = position in B + 1, dlo <= 0 <= dhi) and let every cell outside it be inactive
(pssm = -inf).  Then the cells outside the band are pure gap cells and their values have closed forms in
per-row (region U, c - r > dhi) resp. per-column (region L, c - r < dlo) running maxima of values on the band's
boundary diagonals:

  U:  TH(r) = H(cb,r) + ge - c1*ge,  TS(r) = S(cb,r) + go - c1*ge,  cb = r + dhi, c1 = cb + 1,  T = max(TH, TS)
      M(r)  = max_{r' < r} (T(r') - r'*ge)                       (M(0) = -inf)
      V(c,r) = (c + r - 1)*ge + go + M(r)                         (r >= 1)
      H(c,r) = c*ge + T(r)                 if c == c1 (first cell right of the band)
             = c*ge + max(T(r), K(r))      else,   K(r) = M(r) + 2*go + (r - 2)*ge
  L:  the same with rows and columns, H and V exchanged.

With these, the values and SeqAn trace decisions inside the band equal those of the full matrix, and a traceback
that leaves the band can be continued exactly.

Sheared band (specification for the next kernel): the same closed forms hold when the band's
diagonals move right with the rows, dlo[r] <= c - r <= dhi[r] with both bounds non-decreasing in r and a shift smaller
than the width (dlo[r] <= dhi[r-1]); cb = r + dhi[r] per row, and the L records belong to the lowest band cell of each
column (several columns close in a row where the band shifts).  A band cell that enters on the right takes its upper
and diagonal neighbour from region U's closed forms of the row above.  This file checks the claim against a plain full-matrix
restatement (same tie rules as oracle/lara_oracle.cpp) on random inputs; tests/test_band_proto.py runs it.
Requires go < ge (opening dearer than extending), the reference's normal case.
"""
import numpy as np

NEG = -(1 << 29)          # "minus infinity" of not-yet-existing cells (never wins, never overflows)
PINF = -(1 << 30)         # inactive pssm cell
GAP, FROMH, VOPEN, HOPEN = 1, 2, 4, 8


def cell(D, Hl, Sl, Vu, Su, go, ge):
    """One inner cell.  Returns S, H, V, nibble.  Tie rules: SeqAn _maxScore keeps the first operand."""
    nib = 0
    he, ho = Hl + ge, Sl + go
    if he < ho:
        h = ho; nib |= HOPEN
    else:
        h = he
    ve, vo = Vu + ge, Su + go
    if ve < vo:
        v = vo; nib |= VOPEN
    else:
        v = ve
    if v < h:
        g = h; nib |= FROMH
    else:
        g = v
    if D < g:
        s = g; nib |= GAP
    else:
        s = D
    return s, h, v, nib


def full_dp(la, lb, pssm, go, ge):
    """Plain full matrix (column 0 / row 0 like SeqAn).  Returns S, H, V, nib arrays [(la+1),(lb+1)]."""
    S = np.full((la + 1, lb + 1), NEG, np.int64); H = S.copy(); V = S.copy()
    N = np.zeros((la + 1, lb + 1), np.int64)
    S[0, 0] = 0
    for r in range(1, lb + 1):
        ve, vo = V[0, r - 1] + ge, S[0, r - 1] + go
        V[0, r] = vo if ve < vo else ve
        N[0, r] = VOPEN if ve < vo else 0
        S[0, r] = V[0, r]
    for c in range(1, la + 1):
        he, ho = H[c - 1, 0] + ge, S[c - 1, 0] + go
        H[c, 0] = ho if he < ho else he
        N[c, 0] = HOPEN if he < ho else 0
        S[c, 0] = H[c, 0]
        for r in range(1, lb + 1):
            S[c, r], H[c, r], V[c, r], N[c, r] = cell(S[c - 1, r - 1] + pssm[c - 1, r - 1], H[c - 1, r], S[c - 1, r],
                                                      V[c, r - 1], S[c, r - 1], go, ge)
    return S, H, V, N


def walk(la, lb, nib_at, vEnd, hEnd, sEnd):
    """SeqAn traceback as a state machine reading one cell per transition (see kernels_post.cu).
    nib_at(c, r) -> nibble.  Returns list of lines (a, b) in alignment order."""
    c, r = la, lb
    lines = []
    state = 0
    first = True
    while c > 0 and r > 0:
        nib = nib_at(c, r)
        if state == 0:
            gap, fromh = bool(nib & GAP), bool(nib & FROMH)
            if first:
                first = False
                if vEnd == sEnd:
                    gap, fromh = True, False
                elif hEnd == sEnd:
                    gap, fromh = True, True
            if not gap:
                c -= 1; r -= 1
                lines.append((c, r))
            elif not fromh:
                run = (not (nib & VOPEN)) and r != 1
                r -= 1
                state = 1 if run else 0
            else:
                run = (not (nib & HOPEN)) and c != 1
                c -= 1
                state = 2 if run else 0
        elif state == 1:
            run = (not (nib & VOPEN)) and r != 1
            r -= 1
            state = 1 if run else 0
        else:
            run = (not (nib & HOPEN)) and c != 1
            c -= 1
            state = 2 if run else 0
    return lines[::-1]


class Band:
    """Banded fill in anti-diagonal order with closed-form inflow from U and L, and an exact nib_at for ANY cell."""

    def __init__(self, la, lb, pssm, go, ge, dlo, dhi):
        """dlo, dhi: ints (straight band) or per-row sequences dlo[r], dhi[r] for r = 0..lb, both NON-DECREASING in r
        (sheared band: the diagonals follow the alignment corridor of a pair with la > lb), dlo[0] <= 0 <= dhi[0]."""
        dloR = np.full(lb + 1, dlo, np.int64) if np.isscalar(dlo) else np.asarray(dlo, np.int64)
        dhiR = np.full(lb + 1, dhi, np.int64) if np.isscalar(dhi) else np.asarray(dhi, np.int64)
        assert go < ge and dloR[0] <= 0 <= dhiR[0] and len(dloR) == lb + 1 and len(dhiR) == lb + 1
        assert (np.diff(dloR) >= 0).all() and (np.diff(dhiR) >= 0).all() and (dloR <= dhiR).all()
        # connected: the column left of a row's first band cell has a band cell in the row above (shift < width), so
        # that every column goes U -> band -> L from top to bottom
        assert (dloR[1:] <= dhiR[:-1]).all()
        self.la, self.lb, self.go, self.ge, self.dlo, self.dhi = la, lb, go, ge, dloR, dhiR
        S = {}; H = {}; V = {}; N = {}
        # U arrays per row, L arrays per column; rb[c] = row of the lowest band cell of column c
        self.TH = np.full(lb + 2, NEG, np.int64); self.TS = self.TH.copy(); self.M = np.full(lb + 2, PINF, np.int64)
        self.TV2 = np.full(la + 2, NEG, np.int64); self.TS2 = self.TV2.copy(); self.M2 = np.full(la + 2, PINF, np.int64)
        self.rb = np.full(la + 2, -1, np.int64)
        runM = PINF      # running max for U, over rows already closed
        runM2 = PINF     # running max for L, over columns already closed
        in_band = lambda c, r: 0 <= r <= lb and dloR[r] <= c - r <= dhiR[r]
        for r in range(0, lb + 1):
            for c in range(max(0, r + int(dloR[r])), min(la, r + int(dhiR[r])) + 1):
                d = c - r
                if c == 0 and r == 0:
                    S[(0, 0)] = 0; H[(0, 0)] = NEG; V[(0, 0)] = NEG; N[(0, 0)] = 0
                else:
                    # neighbours: band values, or closed forms of the gap regions.  The upper and the diagonal neighbour
                    # are band or U cells (dlo does not decrease), the left neighbour is a band or an L cell.
                    if r >= 1 and not in_band(c, r - 1):
                        Vu, Su = self.u_V(c, r - 1), self.u_S(c, r - 1)
                    else:
                        Vu, Su = V.get((c, r - 1), NEG), S.get((c, r - 1), NEG)
                    if c >= 1 and not in_band(c - 1, r):
                        Hl, Sl = self.l_H(c - 1, r), self.l_S(c - 1, r)
                    else:
                        Hl, Sl = H.get((c - 1, r), NEG), S.get((c - 1, r), NEG)
                    sub = pssm[c - 1, r - 1] if (c >= 1 and r >= 1) else PINF
                    if c >= 1 and r >= 1 and not in_band(c - 1, r - 1):
                        Sd = self.u_S(c - 1, r - 1)                  # the band moved right: U cell of the row above
                    else:
                        Sd = S.get((c - 1, r - 1), NEG)
                    S[(c, r)], H[(c, r)], V[(c, r)], N[(c, r)] = cell(Sd + sub, Hl, Sl, Vu, Su, go, ge)
                # boundary bookkeeping (after the cell is final)
                if d == dhiR[r]:
                    c1 = c + 1
                    self.M[r] = runM
                    self.TH[r] = H[(c, r)] + ge - c1 * ge
                    self.TS[r] = S[(c, r)] + go - c1 * ge
                    runM = max(runM, max(self.TH[r], self.TS[r]) - r * ge)
                if r == lb or not in_band(c, r + 1):
                    r1 = r + 1                                       # lowest band cell of column c
                    self.rb[c] = r
                    self.M2[c] = runM2
                    self.TV2[c] = V[(c, r)] + ge - r1 * ge
                    self.TS2[c] = S[(c, r)] + go - r1 * ge
                    runM2 = max(runM2, max(self.TV2[c], self.TS2[c]) - c * ge)
        self.S, self.H, self.V, self.N = S, H, V, N

    # ---- region U (c - r > dhi): rows r with r + dhi + 1 <= la ------------------------------------------
    def u_T(self, r): return max(self.TH[r], self.TS[r])
    def u_K(self, r): return self.M[r] + 2 * self.go + (r - 2) * self.ge
    def u_V(self, c, r):
        return NEG if r == 0 else (c + r - 1) * self.ge + self.go + self.M[r]
    def u_H(self, c, r):
        first = (c == r + self.dhi[r] + 1)
        return c * self.ge + (self.u_T(r) if first or r == 0 else max(self.u_T(r), self.u_K(r)))
    def u_S(self, c, r): return max(self.u_V(c, r), self.u_H(c, r))
    def u_nib(self, c, r):
        """Trace nibble of a U cell (r >= 1): diagonal never wins."""
        go, ge = self.go, self.ge
        nib = GAP
        first = (c == r + self.dhi[r] + 1)
        if first:
            if self.TH[r] < self.TS[r]: nib |= HOPEN
        else:
            # H(c-1,r)+ge vs S(c-1,r)+go with S = max(V,H): opens iff V(c-1,r)+go > H(c-1,r)+ge
            if self.u_H(c - 1, r) + ge < self.u_S(c - 1, r) + go: nib |= HOPEN
        if self.u_V(c, r - 1) + ge < self.u_S(c, r - 1) + go: nib |= VOPEN
        if self.u_V(c, r) < self.u_H(c, r): nib |= FROMH
        return nib

    # ---- region L (c - r < dlo) -------------------------------------------------------------------------------
    def l_T(self, c): return max(self.TV2[c], self.TS2[c])
    def l_K(self, c): return self.M2[c] + 2 * self.go + (c - 2) * self.ge
    def l_H(self, c, r):
        return NEG if c == 0 else (c + r - 1) * self.ge + self.go + self.M2[c]
    def l_V(self, c, r):
        first = (r == self.rb[c] + 1)
        return r * self.ge + (self.l_T(c) if first or c == 0 else max(self.l_T(c), self.l_K(c)))
    def l_S(self, c, r): return max(self.l_V(c, r), self.l_H(c, r))
    def l_nib(self, c, r):
        go, ge = self.go, self.ge
        nib = GAP
        first = (r == self.rb[c] + 1)
        if first:
            if self.TV2[c] < self.TS2[c]: nib |= VOPEN
        else:
            if self.l_V(c, r - 1) + ge < self.l_S(c, r - 1) + go: nib |= VOPEN
        if self.l_H(c - 1, r) + ge < self.l_S(c - 1, r) + go: nib |= HOPEN
        if self.l_V(c, r) < self.l_H(c, r): nib |= FROMH
        return nib

    def nib_at(self, c, r):
        d = c - r
        if d > self.dhi[r]: return self.u_nib(c, r)
        if d < self.dlo[r]: return self.l_nib(c, r)
        return self.N[(c, r)]

    def end_values(self):
        c, r = self.la, self.lb
        d = c - r
        if d > self.dhi[r]: return self.u_S(c, r), self.u_V(c, r), self.u_H(c, r)
        if d < self.dlo[r]: return self.l_S(c, r), self.l_V(c, r), self.l_H(c, r)
        return self.S[(c, r)], self.V[(c, r)], self.H[(c, r)]


def check_one(rng, la, lb, dlo, dhi, go, ge, density, alphabet=None):
    """Random pssm with all active cells inside the band; compares band vs full.  Returns None or a message."""
    pssm = np.full((la, lb), PINF, np.int64)
    dloR = np.full(lb + 1, dlo, np.int64) if np.isscalar(dlo) else np.asarray(dlo, np.int64)
    dhiR = np.full(lb + 1, dhi, np.int64) if np.isscalar(dhi) else np.asarray(dhi, np.int64)
    for a in range(la):
        for b in range(lb):
            # cell (c, r) = (a + 1, b + 1) must lie in the band of its row
            if dloR[b + 1] <= a - b <= dhiR[b + 1] and rng.random() < density:
                pssm[a, b] = int(rng.choice(alphabet)) if alphabet is not None else int(rng.integers(-20000, 60000))
    S, H, V, N = full_dp(la, lb, pssm, go, ge)
    band = Band(la, lb, pssm, go, ge, dlo, dhi)
    # 1. values and nibbles inside the band
    for (c, r), s in band.S.items():
        if c == 0 and r == 0:
            continue
        if s != S[c, r] or (c >= 1 and r >= 1 and band.N[(c, r)] != N[c, r]):
            return "band cell (%d,%d): S %d vs %d, nib %d vs %d" % (c, r, s, S[c, r], band.N[(c, r)], N[c, r])
        if c >= 1 and r >= 1 and (band.H[(c, r)] != H[c, r] or band.V[(c, r)] != V[c, r]):
            return "band cell (%d,%d): H/V differ" % (c, r)
    # 2. closed forms outside the band (inner cells)
    for c in range(1, la + 1):
        for r in range(1, lb + 1):
            d = c - r
            if d > dhiR[r]:
                got = (band.u_S(c, r), band.u_H(c, r), band.u_V(c, r), band.u_nib(c, r))
            elif d < dloR[r]:
                got = (band.l_S(c, r), band.l_H(c, r), band.l_V(c, r), band.l_nib(c, r))
            else:
                continue
            want = (S[c, r], H[c, r], V[c, r], N[c, r])
            if got != want:
                return "outside cell (%d,%d) d=%d: got %r want %r" % (c, r, d, got, want)
    # 3. traceback
    sE, vE, hE = band.end_values()
    if (sE, vE, hE) != (S[la, lb], V[la, lb], H[la, lb]):
        return "end values differ"
    l1 = walk(la, lb, lambda c, r: N[c, r], V[la, lb], H[la, lb], S[la, lb])
    l2 = walk(la, lb, band.nib_at, vE, hE, sE)
    if l1 != l2:
        return "lines differ"
    return None


def sheared_bounds(rng, la, lb, width):
    """A band of `width` diagonals that follows the corridor from (0, 0) to (la, lb) of a pair with la >= lb: lower bound
    dlo[r] non-decreasing from <= 0 to about la - lb - width / 2, in random steps of 0..3 diagonals."""
    dlo = np.zeros(lb + 1, np.int64)
    start = -int(rng.integers(0, max(1, width // 2 + 1)))
    target = la - lb - int(rng.integers(0, width))
    dlo[0] = start
    for r in range(1, lb + 1):
        want = start + (target - start) * r // max(lb, 1)
        step = int(rng.integers(0, 4)) if dlo[r - 1] < want else 0
        dlo[r] = dlo[r - 1] + min(step, width - 1)               # connected band: shift < width
    dhi = np.maximum.accumulate(np.maximum(dlo + width - 1, 0))
    return dlo, dhi


if __name__ == "__main__":
    rng = np.random.default_rng(1)
    n = 0
    for it in range(400):
        la = int(rng.integers(1, 40)); lb = int(rng.integers(1, la + 1))
        dlo = -int(rng.integers(0, 6)); dhi = int(rng.integers(0, 8))
        go, ge = (-49152, -16384) if it % 3 else (-3 * 8192, -8192)
        alpha = None if it % 2 else [-16384, 0, 16384, 32768]
        msg = check_one(rng, la, lb, dlo, dhi, go, ge, float(rng.uniform(0.2, 0.9)), alpha)
        if msg:
            print("FAIL it=%d la=%d lb=%d band=[%d,%d] go=%d: %s" % (it, la, lb, dlo, dhi, go, msg))
            n += 1
            if n > 5:
                break
    print("done, failures:", n)
    n = 0
    for it in range(300):
        lb = int(rng.integers(1, 30)); la = lb + int(rng.integers(0, 40))
        dlo, dhi = sheared_bounds(rng, la, lb, int(rng.integers(2, 10)))
        go, ge = (-49152, -16384) if it % 3 else (-3 * 8192, -8192)
        alpha = None if it % 2 else [-16384, 0, 16384, 32768]
        msg = check_one(rng, la, lb, dlo, dhi, go, ge, float(rng.uniform(0.2, 0.9)), alpha)
        if msg:
            print("FAIL sheared it=%d la=%d lb=%d go=%d: %s" % (it, la, lb, go, msg))
            n += 1
            if n > 5:
                break
    print("sheared done, failures:", n)
