"""Vehicle contacts from the other side.  The reference asks Bullet for the bodies within 0.05 m of the
ego's chassis (smarts/core/smarts.py:1270-1291 -> chassis.py:60-80: getOverlappingObjects + getClosestPoints(distance=0.05));
Bullet is absent, so the pair test is this repo's own: a 2-D separating-axis test on the heading-aligned boxes with 0.05 m
of leeway per axis ("parity unpinned").  Here the oracle's answer is bracketed by plain geometry, written
without separating axes: the exact Euclidean distance d between the two rectangles (0 when they overlap: a corner of
one inside the other, or two edges crossing; else the smallest corner-to-edge distance).
    d <= 0.05 m            ->  the pair must be reported as in contact (what getClosestPoints would return);
    d  > 0.05 * sqrt(2) m  ->  it must not (a per-axis leeway reaches at most that far, diagonally off a corner);
and the band in between, where the two definitions may differ, must stay rare.  CPU only; CUDA == oracle bit for bit."""
import numpy as np

from conftest import load_map, make_cfg, oracle_world


def corners(x, y, h, length, width):
    # heading 0 looks along +y (coordinates.py Heading): forward = (-sin h, cos h), right = (cos h, sin h)
    c, s = np.cos(h), np.sin(h)
    out = []
    for a, b in ((-0.5, -0.5), (0.5, -0.5), (0.5, 0.5), (-0.5, 0.5)):
        out.append((x + a * width * c - b * length * s, y + a * width * s + b * length * c))
    return np.array(out)


def point_in_rect(p, r):
    sign = None
    for i in range(4):
        a, b = r[i], r[(i + 1) % 4]
        cr = (b[0] - a[0]) * (p[1] - a[1]) - (b[1] - a[1]) * (p[0] - a[0])
        if abs(cr) < 1e-15:
            continue
        if sign is None:
            sign = cr > 0
        elif (cr > 0) != sign:
            return False
    return True


def seg_seg_cross(a, b, c, d):
    def orient(p, q, r):
        return (q[0] - p[0]) * (r[1] - p[1]) - (q[1] - p[1]) * (r[0] - p[0])
    return (orient(a, b, c) * orient(a, b, d) < 0) and (orient(c, d, a) * orient(c, d, b) < 0)


def point_seg(p, a, b):
    ab, ap = b - a, p - a
    t = np.clip(np.dot(ap, ab) / max(np.dot(ab, ab), 1e-300), 0.0, 1.0)
    return float(np.linalg.norm(ap - t * ab))


def rect_distance(r1, r2):
    if any(point_in_rect(p, r2) for p in r1) or any(point_in_rect(p, r1) for p in r2):
        return 0.0
    for i in range(4):
        for j in range(4):
            if seg_seg_cross(r1[i], r1[(i + 1) % 4], r2[j], r2[(j + 1) % 4]):
                return 0.0
    d = np.inf
    for ra, rb in ((r1, r2), (r2, r1)):
        for p in ra:
            for j in range(4):
                d = min(d, point_seg(p, rb[j], rb[(j + 1) % 4]))
    return d


def test_contacts_bracketed_by_exact_rectangle_distance():
    rm = load_map("4lane")
    rng = np.random.default_rng(11)
    E, A, M = 40, 1, 12
    cfg = make_cfg(E=E, A=A, M=M, events=True, route_cap=16)
    ora = oracle_world(rm, cfg, threads=4)
    ex, ey = rng.uniform(140, 160, (E, A)), rng.uniform(-80, -40, (E, A))
    eh = rng.uniform(-np.pi, np.pi, (E, A))
    ora.reset(ex, ey, eh, np.zeros((E, A)))
    soc = np.zeros((E, M, 8))
    # social boxes scattered a few metres around the ego, many of them grazing it; one in four pushed to the gap scale
    ang, rad = rng.uniform(0, 2 * np.pi, (E, M)), rng.uniform(0.5, 4.5, (E, M))
    soc[..., 0] = ex + rad * np.cos(ang)
    soc[..., 1] = ey + rad * np.sin(ang)
    soc[..., 2] = rng.uniform(-np.pi, np.pi, (E, M))
    soc[..., 4] = rng.choice([3.68, 7.0, 2.5], (E, M))
    soc[..., 5] = rng.choice([1.47, 2.25, 1.0], (E, M))
    soc[..., 6] = 1.5
    soc[..., 7] = rng.uniform(size=(E, M)) < 0.9
    ora.set_social(soc)
    ora.observe()
    o, s = ora.obs(), ora.state()
    must = may = band = 0
    for e in range(E):
        ego = corners(s["pos_x"][e, 0], s["pos_y"][e, 0], s["heading"][e, 0], s["length"][e, 0], s["width"][e, 0])
        lo = hi = 0
        first_sure = None
        for j in range(M):
            if not soc[e, j, 7]:
                continue
            v = 1 + j
            other = corners(s["pos_x"][e, v], s["pos_y"][e, v], s["heading"][e, v], s["length"][e, v], s["width"][e, v])
            d = rect_distance(ego, other)
            if d <= 0.05:
                lo += 1
                hi += 1
                first_sure = v if first_sure is None else first_sure
            elif d <= 0.05 * np.sqrt(2) + 1e-9:
                hi += 1
                band += 1
        n = int(o["ev_n_collisions"][e, 0])
        assert lo <= n <= hi, (e, lo, n, hi)
        assert bool(o["ev_events"][e, 0] & 1) == (n > 0)
        if lo == hi and n:
            assert int(o["ev_collidee"][e, 0]) == first_sure        # the lowest colliding vehicle slot
        must += lo
        may += hi
    assert must > 30 and may - must == band and band <= 0.1 * must + 3
