"""Coverage check of the general ring schedule of csrc/eigh_ring.cuh (RingCols::phases): K0 processors hold columns
(P, Q); ring(g): g rounds, Q moves on by one processor inside its group of g; even g: the halves exchange (P's to the
first half, Q's to the second); odd g > 1: terminal round robin - round r: processor i swaps with (2r - i) mod g
(the one with (i - r) mod g > (g-1)/2 gives its Q and takes the partner's P), rotates, swaps back.
Prints rounds per sweep and verifies that every pair of columns meets at least once."""
import itertools
import sys


def sweep(K0):
    P = [2 * k for k in range(K0)]
    Q = [2 * k + 1 for k in range(K0)]
    met = {}
    rounds = 0

    def rotate():
        nonlocal rounds
        rounds += 1
        for k in range(K0):
            key = frozenset((P[k], Q[k]))
            met[key] = met.get(key, 0) + 1

    def ring(g, bases):
        for r in range(g):
            rotate()
            if r + 1 < g:
                newQ = list(Q)
                for k in range(K0):
                    b = bases[k]
                    newQ[k] = Q[b + (k - b - 1) % g]
                Q[:] = newQ

    g = K0
    bases = [0] * K0
    while True:
        ring(g, bases)
        if g == 1:
            break
        if g % 2 == 0:
            half = g // 2
            for k in range(K0):
                if k - bases[k] < half:
                    kp = k + half
                    Q[k], P[kp] = P[kp], Q[k]
            for k in range(K0):
                if k - bases[k] >= half:
                    bases[k] += half
            g = half
        else:
            h = (g - 1) // 2
            for r in range(g):
                def swap():
                    for k in range(K0):
                        i = k - bases[k]
                        if i == r:
                            continue
                        first = ((i - r) % g) > h
                        if first:
                            kp = bases[k] + (2 * r - i) % g
                            Q[k], P[kp] = P[kp], Q[k]
                swap()
                rotate()
                swap()
            break
    n = 2 * K0
    missing = [c for c in itertools.combinations(range(n), 2) if frozenset(c) not in met]
    return rounds, missing, max(met.values())


if __name__ == '__main__':
    for K0 in (1, 2, 3, 4, 5, 6, 7, 8, 10, 12, 14, 16):
        rounds, missing, mx = sweep(K0)
        print("K0 = %2d: %2d columns, %2d rounds (ideal %2d), pairs never met: %d, most meetings of a pair: %d"
              % (K0, 2 * K0, rounds, 2 * K0 - 1, len(missing), mx))
        assert not missing
