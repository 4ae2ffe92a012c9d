"""Min-weight set cover of the GEM stage for the GPU box (TEST INFRASTRUCTURE; my own statement of Algorithm 2 of the paper as
``superdsm/minsetcover.py:4-82`` and ``:111-151`` implement it): greedy cover by price, merge step, retries with a decreasing
sparsity parameter, solved per cluster.  ``tests/test_gem_seam.py`` checks it against the cover the reference's own ``MinSetCover``
produced (recorded in ``tests/golden/gem_*.npz``)."""


def _greedy(items, beta):
    """items: list of (footprint frozenset, energy).  Returns indices of the accepted sets."""
    universe = set().union(*[fp for fp, _ in items])
    remaining = list(range(len(items)))
    accepted = []
    while remaining:
        best, best_price = None, None
        for i in remaining:                                   # first minimum wins, in list order
            price = (items[i][1] + beta) / len(items[i][0] & universe)
            if best is None or price < best_price: best, best_price = i, price
        accepted.append(best)
        universe -= items[best][0]
        remaining = [i for i in remaining if items[i][0] & universe]
    return accepted


def _merge(items, accepted, beta):
    weight = lambda i: items[i][1] + beta
    for cand in sorted([i for i in range(len(items)) if i not in accepted], key=weight):
        inside, ok = [], True
        for a in accepted:
            common = len(items[a][0] & items[cand][0])
            if common == 0: continue
            if common < len(items[a][0]): ok = False; break
            inside.append(a)
        if ok and weight(cand) < sum(weight(a) for a in inside):
            accepted = [a for a in accepted if a not in inside] + [cand]
    return accepted


def solve(items, beta, max_iter=5, gamma=0.8):
    first = _merge(items, _greedy(items, beta), beta)
    if max_iter > 1 and beta > 0:
        second = solve(items, beta * gamma, max_iter - 1, gamma)
        value = lambda sol: sum(items[i][1] for i in sol) + beta * len(sol)
        if value(second) < value(first): return second
    return first


def cover_by_cluster(footprints, energies, atom_cluster, beta, stages):
    """The stage's incremental cover: the family grows call by call (`stages`: number of sets after each ``MinSetCover.update``,
    first the atoms), and a cluster is re-solved whenever sets of it arrive.  Returns the sorted list of accepted footprints."""
    by_cluster, solution = {}, {}
    done = 0
    for end in stages:
        touched = set()
        for k in range(done, end):
            c = int(atom_cluster[min(footprints[k]) - 1])
            by_cluster.setdefault(c, []).append(k)
            touched.add(c)
        for c in touched:
            if done == 0: solution[c] = list(by_cluster[c])              # the atoms are the initial solution
            else:
                items = [(frozenset(footprints[k]), energies[k]) for k in by_cluster[c]]
                solution[c] = [by_cluster[c][i] for i in solve(items, beta)]
        done = end
    return sorted(tuple(sorted(footprints[k])) for sol in solution.values() for k in sol)
