"""numpy statement of the reference's window semantics: what every consumer
must see, epoch by epoch.  Uses numpy's own Generator for the shuffles and the reference's
partner loop restated with numpy, so it is independent of the native RNG."""
import numpy as np


def reference_partner(rng, n, me):
    """reference ddl/shuffle.py:32-79"""
    row, col = np.arange(n, dtype=np.int32), np.arange(n, dtype=np.int32)
    if n == 1:
        return 0, 0
    if n == 2:
        return (1, 1) if me == 0 else (0, 0)
    while True:
        rng.shuffle(col)
        bad = False
        for r in range(n):
            if col[row == r][0] == row[col == r][0] or (row == col).any():
                bad = True
                break
        if not bad:
            return int(col[row == me][0]), int(row[col == me][0])


def expected_epochs(n_instances, P, n, B, epochs, fraction, designed=True):
    """seen[i][e] = ids consumer i receives in epoch e (tail dropped), with the global exchange
    active (designed) or dead as in the unmodified reference (as wired)."""
    G = P + 1
    wins = {(i, k): (i * G + k) * 1_000_000 + np.arange(n, dtype=np.int64) for i in range(n_instances)
            for k in range(1, P + 1)}
    rngs = {(i, k): np.random.default_rng(i * G + k) for i in range(n_instances) for k in range(1, P + 1)}
    prngs = {(i, k): np.random.default_rng(k) for i in range(n_instances) for k in range(1, P + 1)}
    its = {k: 0 for k in range(1, P + 1)}
    ne = int(n * fraction)
    h = ne // 2
    seen = [[] for _ in range(n_instances)]
    for e in range(epochs):
        k = e % P + 1
        while its[k] <= e // P:
            if designed and n_instances > 1 and ne > 0:
                part = [reference_partner(prngs[(i, k)], n_instances, i) for i in range(n_instances)]
                old = {i: wins[(i, k)].copy() for i in range(n_instances)}
                for i in range(n_instances):
                    send_to, recv_from = part[i]
                    wins[(i, k)][:h] = old[recv_from][:h]
                    wins[(i, k)][h:ne] = old[send_to][h:ne]
            for i in range(n_instances):
                rngs[(i, k)].shuffle(wins[(i, k)])
            its[k] += 1
        for i in range(n_instances):
            seen[i].append(wins[(i, k)][: (n // B) * B].copy())
    return [np.stack(s) for s in seen]
