"""CPU property test of the quantised-node slab test (csrc/bvh_build.cu: k_quantize_nodes, csrc/trace.cuh: nodeTest), restated
in numpy with the kernels' operation order (fp32, FMA = one rounding): a ray that passes through a child's TRUE box must be
accepted by the test on the quantised, padded box — for nearby and for very distant origins, for axis-parallel directions, for
large and for tiny nodes far from the world origin. This is the property the hit-id parity rests on: the box test
may accept too much, never too little."""
import numpy as np

F = np.float32


def fma(a, b, c):
    return F(np.float64(a) * np.float64(b) + np.float64(c))


def quantize(lo, hi):
    """k_quantize_nodes for one node: lo / hi [4, 3] fp32 padded child boxes -> (origin[3] f32, scale[3] f32, qlo[4,3], qhi[4,3])."""
    origin, scale = np.zeros(3, F), np.zeros(3, F)
    qlo, qhi = np.zeros((4, 3), np.int64), np.zeros((4, 3), np.int64)
    for a in range(3):
        mn, mx = F(lo[:, a].min()), F(hi[:, a].max())
        e = -100
        ext0 = float(mx) - float(mn)
        if ext0 > 0:
            e = max(-100, min(100, int(np.frexp(ext0 / 254.0)[1])))
        x = np.float64(mn) - 0.04 * 2.0 ** e                     # __double2float_rd: the largest float <= x
        org = F(x)
        if np.float64(org) > x:
            org = np.nextafter(org, F(-np.inf))
        ext = float(mx) - float(org)
        if ext > 0:
            e = max(-100, min(100, int(np.frexp(ext / 254.5)[1])))
        q = 2.0 ** e
        origin[a], scale[a] = org, F(2.0 ** (e + 15))
        for c in range(4):
            qlo[c, a] = int(min(max(np.floor((float(lo[c, a]) - float(org)) / q - 0.01), 0), 255))
            qhi[c, a] = int(min(max(np.ceil((float(hi[c, a]) - float(org)) / q + 0.01), 0), 255))
    return origin, scale, qlo, qhi


def safe_rcp(d):
    d = F(d)
    return F(1.0) / (d if abs(d) > F(1e-20) else F(np.copysign(1e-20, d)))


def accepts(o, d, mint, maxt, origin, scale, qlo, qhi, child):
    """nodeTest for one child."""
    tn, tf = F(mint), F(maxt)
    for a in range(3):
        idir = safe_rcp(d[a])
        oi = F(F(o[a]) * idir)
        S = F(idir * scale[a])
        C = F(fma(origin[a], idir, -oi) - S)
        e = F(abs(oi) * F(4.76837158e-7))
        Cn, Cf = F(C - e), F(C + e)
        neg = idir < 0
        qn, qf = (qhi, qlo) if neg else (qlo, qhi)
        mn = np.array([(0x3F800000 | (int(qn[child, a]) << 8))], np.uint32).view(F)[0]
        mf = np.array([(0x3F800000 | (int(qf[child, a]) << 8))], np.uint32).view(F)[0]
        tn = max(tn, fma(mn, S, Cn))
        tf = min(tf, fma(mf, S, Cf))
    return tn <= tf


def test_quantised_slab_test_never_rejects_a_ray_through_the_true_box():
    rng = np.random.Generator(np.random.PCG64(11))
    worst = 0
    for trial in range(600):
        world = 10.0 ** rng.uniform(-1, 3)                       # scene extent: coordinates within +-world
        pad = F(4e-6 * world)
        node_ext = world * 10.0 ** rng.uniform(-6, 0)            # from tiny nodes (1e-6 of the scene) to the whole scene
        centre = rng.uniform(-world, world, 3) * (1 - node_ext / world / 2)
        true_lo = np.empty((4, 3)); true_hi = np.empty((4, 3))
        for c in range(4):
            a, b = rng.uniform(-0.5, 0.5, 3) * node_ext, rng.uniform(-0.5, 0.5, 3) * node_ext
            flat = rng.random(3) < 0.15                          # flat boxes (axis-aligned triangles)
            b = np.where(flat, a, b)
            true_lo[c], true_hi[c] = centre + np.minimum(a, b), centre + np.maximum(a, b)
        true_lo, true_hi = true_lo.astype(F), true_hi.astype(F)
        lo, hi = (true_lo - pad).astype(F), (true_hi + pad).astype(F)
        origin, scale, qlo, qhi = quantize(lo, hi)
        for _ in range(40):
            c = int(rng.integers(0, 4))
            p = true_lo[c].astype(np.float64) + rng.random(3) * (true_hi[c].astype(np.float64) - true_lo[c].astype(np.float64))
            d = rng.normal(size=3)
            kind = rng.integers(0, 4)
            if kind == 0:
                d[rng.integers(0, 3)] = 0.0                      # axis-parallel in one component
            elif kind == 1:
                ax = rng.integers(0, 3); d[:] = 0.0; d[ax] = rng.choice([-1.0, 1.0])
            d /= np.linalg.norm(d)
            dist = world * 10.0 ** rng.uniform(-4, 2.7)          # up to ~500 scene extents away
            o = (p - d * dist).astype(F)
            d32 = d.astype(F)
            # the fp32 ray passes within |o| * 2^-23-ish of p; the box test must still accept at t ~ dist
            ok_inf = accepts(o, d32, F(0), F(np.inf), origin, scale, qlo, qhi, c)
            ok_cut = accepts(o, d32, F(0), F(dist * (1 + 1e-5) + 1e-30), origin, scale, qlo, qhi, c)
            worst += int(not ok_inf) + int(not ok_cut)
            assert ok_inf and ok_cut, (trial, world, node_ext, dist, c, o, d32)
    assert worst == 0


def test_quantised_boxes_contain_the_padded_boxes_and_empty_slots_are_inverted():
    rng = np.random.Generator(np.random.PCG64(12))
    for _ in range(300):
        s = 10.0 ** rng.uniform(-5, 3)
        c = rng.uniform(-100, 100, 3)
        a, b = rng.uniform(-s, s, (4, 3)), rng.uniform(-s, s, (4, 3))
        lo, hi = (c + np.minimum(a, b)).astype(F), (c + np.maximum(a, b)).astype(F)
        origin, scale, qlo, qhi = quantize(lo, hi)
        quantum = scale.astype(np.float64) / 32768.0
        dlo = origin.astype(np.float64) + qlo * quantum
        dhi = origin.astype(np.float64) + qhi * quantum
        assert (dlo <= lo.astype(np.float64) - 0.009 * quantum).all() and (dhi >= hi.astype(np.float64) + 0.009 * quantum).all()
        assert (qlo >= 0).all() and (qhi <= 255).all() and (qhi > qlo).all()
