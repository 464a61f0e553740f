"""Deterministic synthetic genome sets of the shapes BASELINE.json names.

No sequence data ships with the reference (examples/*.tsv are accession lists
that GetData would download) and there is no network, so every benchmark and
parity input is generated here.  Sequence names use the
reference's ``GENOME&chrN&c`` form (src/model/Sequence.hpp:90-107) so that
``Sequence::genome()`` parses.  Output alphabet is strictly upper-case ATGCN,
i.e. already ``Sequence::to_atgcn``-clean (src/model/Sequence.cpp:151-179).

Model: one ancestral genome per chromosome with i.i.d. bases at the stated GC;
each strain = ancestor + SNPs at rate (1-identity) + indels (1-30 bp) at 1/20
of the SNP rate + inversions and translocations of 5-50 kb; insertion-sequence
families pasted `copies` times per strain at random places and orientations;
runs of N (10-2000 bp).
"""
import numpy as np

_A, _T, _G, _C, _N = (ord(c) for c in "ATGCN")
_COMP = np.arange(256, dtype=np.uint8)
for a, b in ((_A, _T), (_T, _A), (_G, _C), (_C, _G)):
    _COMP[a] = b

CONFIGS = {
    # name: dict(strains, chrs, gc, identity, inv, trans, is_families, n_runs, seed)
    "tiny": dict(strains=3, chrs=[6000, 3500], gc=0.55, identity=0.99, inv=1,
                 trans=1, is_families=[(300, 3)], n_runs=2, seed=7,
                 sv_len=(200, 800)),
    "small": dict(strains=4, chrs=[210000, 119000], gc=0.572, identity=0.995,
                  inv=2, trans=0, is_families=[(842, 5)], n_runs=3, seed=11),
    # C1: AnchorFinder on 3 synthetic Brucella-shaped genomes (CPU config)
    "brucella3": dict(strains=3, chrs=[2110000, 1190000], gc=0.572,
                      identity=0.995, inv=2, trans=0, is_families=[(842, 7)],
                      n_runs=3, seed=1001),
    # C2: 17 synthetic Brucella-shaped genomes (~56 Mbp) on 1 B200
    "brucella17": dict(strains=17, chrs=[2110000, 1190000], gc=0.572,
                       identity=0.995, inv=2, trans=0, is_families=[(842, 7)],
                       n_runs=3, seed=1002),
    # C3: Burkholderia cenocepacia-shaped, 3 chr, high GC
    "burkholderia": dict(strains=5, chrs=[3870000, 3220000, 880000], gc=0.668,
                         identity=0.98, inv=6, trans=4,
                         is_families=[(1200, 10), (1350, 10), (1500, 10)],
                         n_runs=3, seed=1003),
    # C4: Yersinia pestis-shaped, dense IS-element repeats
    "yersinia": dict(strains=12, chrs=[4650000], gc=0.476, identity=0.999,
                     inv=20, trans=0,
                     is_families=[(1954, 110), (708, 65), (1322, 25),
                                  (1100, 10), (1100, 10)],
                     n_runs=5, seed=1004),
    # C5: 1 Gbp pangenome: 30 clusters x 10 strains x (2.1 + 1.23 Mbp)
    "pangenome1g": dict(clusters=30, strains=10, chrs=[2100000, 1230000],
                        gc=(0.35, 0.68), identity=0.99, inv=2, trans=0,
                        is_families=[(1000, 10), (1400, 10)], n_runs=2,
                        seed=1005),
}


def _random_bases(rng, n, gc):
    """i.i.d. bases with P(G)+P(C) = gc."""
    u = rng.random(n, dtype=np.float32)
    half_at = (1.0 - gc) / 2.0
    out = np.full(n, _C, dtype=np.uint8)
    out[u < half_at] = _A
    out[(u >= half_at) & (u < 2 * half_at)] = _T
    out[(u >= 2 * half_at) & (u < 2 * half_at + gc / 2.0)] = _G
    return out


def revcomp(a):
    return _COMP[a[::-1]]


def _mutate(rng, anc, identity, gc):
    seq = anc.copy()
    n = len(seq)
    rate = 1.0 - identity
    # SNPs: replace by a different base
    idx = np.flatnonzero(rng.random(n, dtype=np.float32) < rate)
    if len(idx):
        alphabet = np.array([_A, _T, _G, _C], dtype=np.uint8)
        new = alphabet[rng.integers(0, 4, len(idx))]
        same = new == seq[idx]
        new[same] = _COMP[new[same]]
        seq[idx] = new
    # indels, 1/20 of the SNP rate, 1..30 bp, half deletions
    n_indel = int(rng.binomial(n, rate / 20.0)) if rate > 0 else 0
    if n_indel:
        pos = np.sort(rng.integers(0, n, n_indel))
        lens = rng.integers(1, 31, n_indel)
        is_del = rng.random(n_indel) < 0.5
        keep = np.ones(n, dtype=bool)
        for p, l in zip(pos[is_del], lens[is_del]):
            keep[p:p + l] = False
        ins_pos = pos[~is_del]
        ins_len = lens[~is_del]
        if len(ins_pos):
            ins_bases = _random_bases(rng, int(ins_len.sum()), gc)
            where = np.repeat(ins_pos, ins_len)
            # positions refer to the un-deleted coordinate system
            keep_ins = np.insert(keep, where, True)
            seq = np.insert(seq, where, ins_bases)
            seq = seq[keep_ins]
        else:
            seq = seq[keep]
    return seq


def _structural(rng, seq, n_inv, n_trans, sv_len):
    lo, hi = sv_len
    for _ in range(n_inv):
        l = int(rng.integers(lo, hi + 1))
        if l >= len(seq):
            continue
        s = int(rng.integers(0, len(seq) - l))
        seq[s:s + l] = revcomp(seq[s:s + l])
    for _ in range(n_trans):
        l = int(rng.integers(lo, hi + 1))
        if l >= len(seq):
            continue
        s = int(rng.integers(0, len(seq) - l))
        piece = seq[s:s + l].copy()
        rest = np.concatenate([seq[:s], seq[s + l:]])
        d = int(rng.integers(0, len(rest) + 1))
        seq = np.concatenate([rest[:d], piece, rest[d:]])
    return seq


def _strain(rng, ancestors, cfg, gc, elements):
    sv_len = cfg.get("sv_len", (5000, 50000))
    chrs = [_structural(rng, _mutate(rng, a, cfg["identity"], gc),
                        cfg["inv"], cfg["trans"], sv_len) for a in ancestors]
    # IS elements: `copies` per strain over all chromosomes
    for elem, copies in elements:
        for _ in range(copies):
            c = int(rng.integers(0, len(chrs)))
            p = int(rng.integers(0, len(chrs[c]) + 1))
            e = elem if rng.random() < 0.5 else revcomp(elem)
            chrs[c] = np.insert(chrs[c], np.full(len(e), p), e)
    # runs of N spread over the strain
    for _ in range(cfg["n_runs"]):
        c = int(rng.integers(0, len(chrs)))
        l = int(rng.integers(10, 2001))
        l = min(l, max(1, len(chrs[c]) // 4))
        p = int(rng.integers(0, len(chrs[c]) - l + 1))
        chrs[c][p:p + l] = _N
    return chrs


def generate(config, scale=1.0, seed=None):
    """Returns (names, seqs): list of str names, list of uint8 ASCII arrays.

    `scale` shrinks every chromosome length (bounded CPU samples keep the
    workload's shape: same strains, identity and repeat structure)."""
    cfg = dict(CONFIGS[config]) if isinstance(config, str) else dict(config)
    rng = np.random.Generator(np.random.PCG64(cfg["seed"] if seed is None else seed))
    chr_lens = [max(64, int(l * scale)) for l in cfg["chrs"]]
    clusters = cfg.get("clusters", 1)
    names, seqs = [], []
    for cl in range(clusters):
        gc = cfg["gc"]
        if isinstance(gc, tuple):
            gc = float(rng.uniform(gc[0], gc[1]))
        ancestors = [_random_bases(rng, l, gc) for l in chr_lens]
        elements = [(_random_bases(rng, min(l, max(20, min(chr_lens) // 8)), gc), c)
                    for (l, c) in cfg["is_families"]]
        for st in range(cfg["strains"]):
            chrs = _strain(rng, ancestors, cfg, gc, elements)
            g = "G%02dS%02d" % (cl, st) if clusters > 1 else "STRAIN%02d" % st
            for ci, c in enumerate(chrs):
                names.append("%s&chr%d&c" % (g, ci + 1))
                seqs.append(np.ascontiguousarray(c))
    return names, seqs


def generate_shard(config, rank, world, scale=1.0):
    """Weak-scaling input of rank `rank`: its own cluster of the job (same
    shape as `config`, an independent ancestor per rank).  Returns (names,
    seqs); names carry the rank so that genomes stay distinct job-wide."""
    cfg = dict(CONFIGS[config]) if isinstance(config, str) else dict(config)
    names, seqs = generate(cfg, scale=scale, seed=cfg["seed"] + 7919 * rank)
    return ["R%02d%s" % (rank, n) for n in names], seqs


def genome_ids(names):
    """Integer labels for Sequence::genome() (src/model/Sequence.cpp:193-202)."""
    ids, table = [], {}
    for n in names:
        parts = n.split("&")
        g = parts[0] if len(parts) == 3 and parts[2] in ("c", "l") else ""
        ids.append(-1 if g == "" else table.setdefault(g, len(table)))
    return ids


def total_bases(seqs):
    return int(sum(len(s) for s in seqs))
