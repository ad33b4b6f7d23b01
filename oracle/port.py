"""TEST INFRASTRUCTURE — numpy restatement of the reference's hot-path arithmetic (fp32, one rounding per
operation, the reference's operation order). It exists for boxes where the compiled reference
(oracle/_ref) is absent, and as an independent cross-check of it; tests/test_port_oracle.py pins it to the
golden vectors that oracle/_ref produced. Nothing in the product imports this file.

Covered: TriAccel::load + rayIntersect (include/mitsuba/render/triaccel.h:63-159), the kd-tree's scene-box
clip and adaptive epsilon (src/librender/skdtree.cpp:108-132,180-199, include/mitsuba/core/aabb.h:349-382),
closest / any hit by exhaustive search with the reference's acceptance rules ("later equal t wins" has no
analogue without the kd-tree's leaf order, so exact ties are reported as the lowest index), BSDF f/pdf of
the five models (src/bsdfs/*.cpp, src/libcore/util.cpp:680-725), the perspective camera
(src/cameras/perspective.cpp:77-112) and ImageBlock::putSample (include/mitsuba/render/imageblock.h:80-138).
"""
import numpy as np

F = np.float32
EPS = F(1e-4)
INV_PI = F(0.31830988618379067154)
PI = F(3.14159265358979323846)


# ---- TriAccel ----------------------------------------------------------------------------------------
def triaccel_load(A, B, C):
    """triaccel.h:63-96 for arrays of triangles [n,3] (fp32). Returns dict of fields; k = 3 when degenerate."""
    A, B, C = (np.asarray(x, F) for x in (A, B, C))
    b, c = C - A, B - A
    N = np.stack([c[:, 1] * b[:, 2] - c[:, 2] * b[:, 1], c[:, 2] * b[:, 0] - c[:, 0] * b[:, 2],
                  c[:, 0] * b[:, 1] - c[:, 1] * b[:, 0]], 1).astype(F)
    k = np.zeros(len(A), np.int64)
    for j in range(3):
        k = np.where(np.abs(N[np.arange(len(A)), j]) > np.abs(N[np.arange(len(A)), k]), j, k)
    u, v = (k + 1) % 3, (k + 2) % 3
    r = np.arange(len(A))
    n_k = N[r, k]
    denom = (b[r, u] * c[r, v] - b[r, v] * c[r, u]).astype(F)
    with np.errstate(divide="ignore", invalid="ignore"):
        dotAN = ((A[:, 0] * N[:, 0] + A[:, 1] * N[:, 1]).astype(F) + A[:, 2] * N[:, 2]).astype(F)
        out = dict(k=np.where(denom == 0, 3, k).astype(np.uint32), n_u=(N[r, u] / n_k).astype(F), n_v=(N[r, v] / n_k).astype(F),
                   n_d=(dotAN / n_k).astype(F), a_u=A[r, u], a_v=A[r, v], b_nu=(b[r, u] / denom).astype(F),
                   b_nv=(-b[r, v] / denom).astype(F), c_nu=(c[r, v] / denom).astype(F), c_nv=(-c[r, u] / denom).astype(F))
    return out


def scene_box(positions_list):
    """Tight box over all vertices, then the two enlarging statements of gkdtree.h:1172-1176."""
    lo = np.min([p.min(0) for p in positions_list], 0).astype(F)
    hi = np.max([p.max(0) for p in positions_list], 0).astype(F)
    nmin = (lo - (((hi - lo).astype(F) * EPS).astype(F) + EPS).astype(F)).astype(F)
    nmax = (hi + (((hi - nmin).astype(F) * EPS).astype(F) + EPS).astype(F)).astype(F)
    return nmin, nmax


def clip_to_scene(ray, box, shadow):
    """AABB::rayIntersect + the interval set-up of ShapeKDTree::rayIntersect. ray: 8 floats. -> (ok, mint, maxt)"""
    o, rmint, d, rmaxt = ray[0:3], ray[3], ray[4:7], ray[7]
    near, far = F(-np.inf), F(np.inf)
    for i in range(3):
        if d[i] == 0:
            if o[i] < box[0][i] or o[i] > box[1][i]:
                return False, F(0), F(0)
        else:
            with np.errstate(all="ignore"):
                rcp = F(1.0) / d[i]
                t1, t2 = F((box[0][i] - o[i]) * rcp), F((box[1][i] - o[i]) * rcp)
            if t1 > t2:
                t1, t2 = t2, t1
            near = max(near, t1); far = min(far, t2)
            if near > far:
                return False, F(0), F(0)
    m = rmint
    if m == EPS:
        s = max(abs(o[0]), abs(o[1]), abs(o[2]))
        if not shadow:
            s = max(s, EPS)
        m = F(m * s)
    mint = m if m > near else near
    maxt = rmaxt if rmaxt < far else far
    return bool(maxt > mint), F(mint), F(maxt)


def trace(ta, occluder, box, rays, shadow=False):
    """Exhaustive closest / any hit with TriAccel::rayIntersect (triaccel.h:98-159) over all triangles.
    ta: dict from triaccel_load (+ optional shape/prim arrays); occluder: bool per triangle.
    closest -> (t, u, v, index) arrays, index -1 on miss; shadow -> uint32 occluded flags."""
    rays = np.asarray(rays, F).reshape(-1, 8)
    n = len(rays)
    k = ta["k"].astype(np.int64)
    valid = k < 3
    ku, kv = (k + 1) % 3, (k + 2) % 3
    kk = np.where(valid, k, 0)
    res_t = np.full(n, np.inf, F); res_u = np.zeros(n, F); res_v = np.zeros(n, F); res_i = np.full(n, -1, np.int64)
    occ = np.zeros(n, np.uint32)
    for i in range(n):
        ok, mint, maxt = clip_to_scene(rays[i], box, shadow)
        if not ok:
            continue
        o, d = rays[i, 0:3], rays[i, 4:7]
        o_u, o_v, o_k, d_u, d_v, d_k = o[ku], o[kv], o[kk], d[ku], d[kv], d[kk]
        with np.errstate(all="ignore"):
            recip = (F(1.0) / ((d_u * ta["n_u"] + d_v * ta["n_v"]).astype(F) + d_k).astype(F)).astype(F)
            t = ((((ta["n_d"] - o_u * ta["n_u"]).astype(F) - o_v * ta["n_v"]).astype(F) - o_k).astype(F) * recip).astype(F)
            hu = ((o_u + t * d_u).astype(F) - ta["a_u"]).astype(F)
            hv = ((o_v + t * d_v).astype(F) - ta["a_v"]).astype(F)
            u = (hv * ta["b_nu"] + hu * ta["b_nv"]).astype(F)
            v = (hu * ta["c_nu"] + hv * ta["c_nv"]).astype(F)
            hit = valid & ~((t < mint) | (t > maxt)) & (u >= 0) & (v >= 0) & ((u + v).astype(F) <= 1.0)
        if shadow:
            occ[i] = 1 if (hit & occluder).any() else 0
        elif hit.any():
            j = int(np.argmin(np.where(hit, t, np.inf)))
            res_t[i], res_u[i], res_v[i], res_i[i] = t[j], u[j], v[j], j
    return occ if shadow else (res_t, res_u, res_v, res_i)


# ---- BSDF f / pdf (full type mask, ERadiance) ------------------------------------------------------------
def _fresnel_dielectric(c1, c2, etaI, etaT):
    Rs = F(F(F(etaI * c1) - F(etaT * c2)) / F(F(etaI * c1) + F(etaT * c2)))
    Rp = F(F(F(etaT * c1) - F(etaI * c2)) / F(F(etaT * c1) + F(etaI * c2)))
    return F(F(F(Rs * Rs) + F(Rp * Rp)) / F(2.0))


def _fresnel(cosI, ext, int_):
    etaI, etaT = (F(ext), F(int_)) if cosI >= 0 else (F(int_), F(ext))
    sinT = F(F(etaI / etaT) * np.sqrt(max(F(0), F(F(1.0) - F(cosI * cosI)))).astype(F))
    if sinT > 1.0:
        return F(1.0)
    cosT = np.sqrt(F(F(1.0) - F(sinT * sinT))).astype(F)
    return _fresnel_dielectric(abs(cosI), cosT, etaI, etaT)


def _tan_theta(v):
    temp = F(F(1) - F(v[2] * v[2]))
    return F(0) if temp <= 0 else F(np.sqrt(temp).astype(F) / v[2])


def _beckmann_d(m, alpha):
    ex = F(_tan_theta(m) / alpha)
    return F(F(np.exp(-F(ex * ex), dtype=F)) / F(F(F(PI * alpha) * alpha) * F(np.power(m[2], F(4.0), dtype=F))))


def _g1(v, m, alpha):
    if F(F(np.dot(v, m)) * v[2]) <= 0:
        return F(0)
    tt = _tan_theta(v)
    if tt == 0:
        return F(1)
    a = F(F(1.0) / F(alpha * tt))
    if a >= 1.6:
        return F(1)
    a2 = F(a * a)
    return F(F(F(F(3.535) * a) + F(F(2.181) * a2)) / F(F(F(1.0) + F(F(2.276) * a)) + F(F(2.577) * a2)))


def _dot(a, b):
    return F(F(F(a[0] * b[0]) + F(a[1] * b[1])) + F(a[2] * b[2]))


def _normalize(v):
    r = F(F(1.0) / np.sqrt(_dot(v, v)).astype(F))
    return (v * r).astype(F)


def _rg_eval_d(m, alpha, distr):                      # roughglass.cpp:209-252
    if m[2] <= 0:
        return F(0)
    with np.errstate(all="ignore"):
        if distr == 0:
            r = _beckmann_d(m, alpha)
        elif distr == 1:
            r = F(F(F(alpha + F(2)) * F(0.15915494309189533577)) * F(np.power(m[2], alpha, dtype=F)))
        else:
            tt, ct = _tan_theta(m), m[2]
            root = F(alpha / F(F(ct * ct) * F(F(alpha * alpha) + F(tt * tt))))
            r = F(INV_PI * F(root * root))
    return F(0) if float(r) < 1e-40 else r


def _rg_g1(v, m, alpha, distr):                        # roughglass.cpp:298-340
    tt = abs(_tan_theta(v))
    if tt == 0:
        return F(1)
    if F(_dot(v, m) * v[2]) <= 0:
        return F(0)
    if distr == 2:
        root = F(alpha * tt)
        return F(F(2.0) / F(F(1.0) + np.sqrt(F(F(1.0) + F(root * root))).astype(F)))
    if distr == 1:
        alpha = F(np.sqrt(F(F(F(0.5) * alpha) + F(1))).astype(F) / tt)
    a = F(F(1.0) / F(alpha * tt))
    if a >= 1.6:
        return F(1)
    a2 = F(a * a)
    return F(F(F(F(3.535) * a) + F(F(2.181) * a2)) / F(F(F(1.0) + F(F(2.276) * a)) + F(F(2.577) * a2)))


def _roughglass_f_pdf(p, wi, wo):
    intI, extI, distr = p[7], p[8], int(p[9])
    alpha = F(F(F(F(F(0) + p[6]) + p[6]) + p[6]) * F(F(1.0) / F(3)))
    reflect = F(wi[2] * wo[2]) > 0
    etaI, etaT = (extI, intI) if wi[2] >= 0 else (intI, extI)
    if reflect:
        H = (_normalize((wo + wi).astype(F)) * (F(-1) if wo[2] < 0 else F(1))).astype(F)
    else:
        H = (_normalize(((wi * etaI).astype(F) + (wo * etaT).astype(F)).astype(F)) * (F(1) if extI > intI else F(-1))).astype(F)
    wiH, woH = _dot(wi, H), _dot(wo, H)
    sqrt_denom = F(F(etaI * wiH) + F(etaT * woH))
    with np.errstate(all="ignore"):
        # ---- f
        D = _rg_eval_d(H, alpha, distr)
        if D == 0:
            f = np.zeros(3, F)
        else:
            Fr = _fresnel(wiH, extI, intI)
            G = F(_rg_g1(wi, H, alpha, distr) * _rg_g1(wo, H, alpha, distr))
            if reflect:
                value = F(F(F(Fr * D) * G) / F(F(F(4.0) * wi[2]) * wo[2]))
                f = (p[0:3] * value).astype(F)
            else:
                num = F(F(F(F(F(F(F(1) - Fr) * D) * G) * etaT) * etaT) * wiH) * woH
                value = F(F(num) / F(F(F(wi[2] * wo[2]) * sqrt_denom) * sqrt_denom))
                value = F(value * F(F(etaI * etaI) / F(etaT * etaT)))
                f = (p[3:6] * abs(value)).astype(F)
        # ---- pdf
        if reflect:
            dwh = F(F(1.0) / F(F(4.0) * woH))
        else:
            dwh = F(F(F(etaT * etaT) * woH) / F(sqrt_denom * sqrt_denom))
        sa = F(alpha * F(F(1.2) - F(F(0.2) * np.sqrt(abs(wi[2])).astype(F))))
        prob = _rg_eval_d(H, sa, distr)
        Fc = min(F(0.9), max(F(0.1), _fresnel(wi[2], extI, intI)))
        prob = F(prob * (Fc if reflect else F(F(1) - Fc)))
        pdf = abs(F(F(prob * H[2]) * dwh))
    return f, F(pdf)


def bsdf_f_pdf(mtype, p, wi, wo):
    """f(wi,wo) [3] and pdf(wi,wo) of one material (type + 16-float block, include/mtsgpu.h layout)."""
    wi, wo, p = np.asarray(wi, F).copy(), np.asarray(wo, F).copy(), np.asarray(p, F)
    zero = np.zeros(3, F)
    if mtype & 0x200:   # mask (mask.cpp:68-82): nested f and pdf scaled by the opacity's luminance
        o = F(F(F(p[13] * F(0.212671)) + F(p[14] * F(0.715160))) + F(p[15] * F(0.072169)))
        f, pdf = bsdf_f_pdf(mtype & ~0x200, p, wi, wo)
        return (f * o).astype(F), F(pdf * o)
    if mtype & 0x100:   # twosided adapter (twosided.cpp:81-100)
        if wi[2] < 0:
            wi[2] *= -1; wo[2] *= -1
        mtype &= 0xFF
    if mtype == 0:      # lambertian.cpp:95-108
        if wi[2] <= 0 or wo[2] <= 0:
            return zero, F(0)
        return (p[0:3] * INV_PI).astype(F), F(wo[2] * INV_PI)
    if mtype == 8:      # difftrans.cpp:91-102
        if F(wi[2] * wo[2]) >= 0:
            return zero, F(0)
        return (p[0:3] * INV_PI).astype(F), F(abs(wo[2]) * INV_PI)
    if mtype == 5:      # phong.cpp:104-162
        if wi[2] <= 0 or wo[2] <= 0:
            return zero, F(0)
        kd, ks, ex, ssw, dsw = p[6], p[7], p[8], p[9], p[10]
        alpha = _dot(np.array([-wi[0], -wi[1], wi[2]], F), wo)
        with np.errstate(all="ignore"):
            pw = F(np.power(alpha, ex, dtype=F))
        spec = F(0) if alpha <= 0 else F(F(F(F(ex + F(2)) * F(0.15915494309189533577)) * pw) * ks)
        f = ((p[3:6] * spec).astype(F) + (p[0:3] * F(INV_PI * kd)).astype(F)).astype(F)
        sp = F(0) if alpha <= 0 else F(F(pw * F(ex + F(1.0))) / F(F(2.0) * PI))
        return f, F(F(ssw * sp) + F(dsw * F(wo[2] * INV_PI)))
    if mtype == 7:      # ward.cpp:145-227
        if wi[2] <= 0 or wo[2] <= 0:
            return zero, F(0)
        kd, ks, ax, ay, ssw, dsw, model = p[6], p[7], p[8], p[9], p[10], p[11], int(p[12])
        H = (wi + wo).astype(F)
        with np.errstate(all="ignore"):
            if model == 0:
                f1 = F(F(1.0) / F(F(F(F(F(4.0) * PI) * ax) * ay) * np.sqrt(F(wi[2] * wo[2])).astype(F)))
            elif model == 1:
                f1 = F(F(1.0) / F(F(F(F(F(F(4.0) * PI) * ax) * ay) * wi[2]) * wo[2]))
            else:       # std::pow(Float, int) is a double; the expression stays double until the assignment
                f1 = F(float(_dot(H, H)) / (float(F(F(PI * ax) * ay)) * float(H[2]) ** 4))
            f2, f3 = F(H[0] / ax), F(H[1] / ay)
            ex = F(-F(F(f2 * f2) + F(f3 * f3)) / F(H[2] * H[2]))
            spec = F(F(f1 * F(np.exp(ex, dtype=F))) * ks)
            f = np.zeros(3, F)
            if spec > F(1e-10):
                f = (f + (p[3:6] * spec).astype(F)).astype(F)
            f = (f + (p[0:3] * F(INV_PI * kd)).astype(F)).astype(F)
            Hn = _normalize(H)
            g1 = F(1.0 / (float(F(F(F(F(F(4.0) * PI) * ax) * ay) * _dot(Hn, wi))) * float(Hn[2]) ** 3))
            g2, g3 = F(Hn[0] / ax), F(Hn[1] / ay)
            gex = F(-F(F(g2 * g2) + F(g3 * g3)) / F(Hn[2] * Hn[2]))
            pspec = F(g1 * F(np.exp(gex, dtype=F)))
        return f, F(F(ssw * pspec) + F(dsw * F(wo[2] * INV_PI)))
    if mtype == 6:      # roughglass.cpp:345-488
        return _roughglass_f_pdf(p, wi, wo)
    if mtype in (2, 4):  # mirror / dielectric: delta only
        return zero, F(0)
    if wi[2] <= 0 or wo[2] <= 0:
        return zero, F(0)
    Hr = _normalize((wi + wo).astype(F))
    if mtype == 1:      # microfacet.cpp:135-199
        kd, ks, alpha, intI, extI = p[6], p[7], p[8], p[9], p[10]
        Fr = _fresnel(_dot(wi, Hr), extI, intI)
        D = _beckmann_d(Hr, alpha)
        G = F(_g1(wi, Hr, alpha) * _g1(wo, Hr, alpha))
        spec = F(F(D * G) / F(F(F(4.0) * wi[2]) * wo[2]))
        f = ((p[3:6] * spec).astype(F) * F(Fr * ks)).astype(F) + (p[0:3] * F(F(INV_PI * F(1 - Fr)) * kd)).astype(F)
        fr = min(max(_fresnel(wi[2], extI, intI), F(0.05)), F(0.95))
        dw, sw = F(F(1 - fr) * kd), F(fr * ks)
        norm = F(F(1) / F(dw + sw))
        pspec = F(F(D * Hr[2]) / F(F(4.0) * abs(_dot(wo, Hr))))
        pdf = F(F(F(sw * pspec) + F(dw * F(wo[2] * INV_PI))) * norm)
        return f.astype(F), pdf
    # roughmetal.cpp:116-145
    alpha, eta, kk = p[3], p[4:7], p[7:10]
    c = _dot(wi, Hr)
    tmpF = (eta * eta + kk * kk).astype(F)
    tmp = (tmpF * F(c * c)).astype(F)
    two = (eta * F(F(2.0) * c)).astype(F)
    rpar = (((tmp - two).astype(F) + F(1)).astype(F) / ((tmp + two).astype(F) + F(1)).astype(F)).astype(F)
    rper = (((tmpF - two).astype(F) + F(c * c)).astype(F) / ((tmpF + two).astype(F) + F(c * c)).astype(F)).astype(F)
    Fc = ((rpar + rper).astype(F) / F(2.0)).astype(F)
    D = _beckmann_d(Hr, alpha)
    G = F(_g1(wi, Hr, alpha) * _g1(wo, Hr, alpha))
    f = (p[0:3] * (Fc * F(F(D * G) / F(F(F(4.0) * wi[2]) * wo[2]))).astype(F)).astype(F)
    pdf = F(F(D * Hr[2]) * F(F(1.0) / F(F(4.0) * abs(_dot(wo, Hr)))))
    return f, pdf


# ---- camera + film -------------------------------------------------------------------------------------
def camera_rays(r2c, c2w, near, far, samples):
    """Pinhole branch of PerspectiveCameraImpl::generateRay (perspective.cpp:77-112), vectorised."""
    r2c, c2w = np.asarray(r2c, F), np.asarray(c2w, F)
    s = np.asarray(samples, F)
    x, y = s[:, 0], s[:, 1]
    p = [((r2c[i, 0] * x + r2c[i, 1] * y).astype(F) + F(r2c[i, 2] * 0)).astype(F) + r2c[i, 3] for i in range(4)]
    w = p[3].astype(F)
    rw = (F(1) / w).astype(F)
    d = np.stack([np.where(w != 1, (p[i] * rw).astype(F), p[i]) for i in range(3)], 1).astype(F)
    ln = np.sqrt(((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]).astype(F) + d[:, 2] * d[:, 2]).astype(F)).astype(F)
    d = (d * (F(1) / ln)[:, None]).astype(F)
    inv_z = (F(1) / d[:, 2]).astype(F)
    out = np.empty((len(s), 8), F)
    out[:, 0:3] = c2w[:3, 3]
    out[:, 3] = F(near) * inv_z; out[:, 7] = F(far) * inv_z
    for i in range(3):
        out[:, 4 + i] = ((c2w[i, 0] * d[:, 0] + c2w[i, 1] * d[:, 1]).astype(F) + c2w[i, 2] * d[:, 2]).astype(F)
    return out


def film_splat(width, height, crop, table, size, samples):
    """ImageBlock::putSample into one full-frame block + Film::putImageBlock (imageblock.h:80-138)."""
    cx, cy, cw, ch = crop
    size = F(size); factor = F(F(15) / size)
    border = int(np.ceil(size - F(0.5)))
    sx, sy = cw + 2 * border, ch + 2 * border
    acc = np.zeros((sy, sx, 5), np.float64)
    for s in np.asarray(samples, F):
        spec = s[2:5]
        if np.isnan(spec).any() or (spec < 0).any():
            continue
        fx = F(F(s[0] - F(0.5)) - F(cx - border)); fy = F(F(s[1] - F(0.5)) - F(cy - border))
        x0, x1 = max(0, int(np.ceil(fx - size))), min(int(np.floor(fx + size)), sx - 1)
        y0, y1 = max(0, int(np.ceil(fy - size))), min(int(np.floor(fy + size)), sy - 1)
        for yy in range(y0, y1 + 1):
            iy = min(int(F(factor * abs(F(yy) - fy))), 15)
            for xx in range(x0, x1 + 1):
                ix = min(int(F(factor * abs(F(xx) - fx))), 15)
                wgt = float(table[iy, ix])
                acc[yy, xx, 0:3] += spec.astype(np.float64) * wgt
                acc[yy, xx, 3] += float(s[5]) * wgt
                acc[yy, xx, 4] += wgt
    return acc[border:border + ch, border:border + cw].astype(F)
