"""CPU simulation of the warp-autonomous ring walk of csrc/blend_lq.cu on the bench workload (MEASUREMENT
INFRASTRUCTURE, not product code; uses the CPU oracle for the projected records / sorted lists / last ids).

For every half tile (one warp: 32 lanes = 32 quads of 2x2 pixels) it builds the per-quad hit masks exactly as
publish_hits does, then replays the kernel's iteration rule -- one hit per lane per iteration, one ring word
retired per iteration at most, a lane steps over an empty word in one iteration -- and counts warp iterations and
live lanes for design variants:

  nw        ring words (32 staged entries each) a warp can hold
  compact   stage only entries whose footprint touches the warp's half tile
  skip      a lane jumps straight to its next non-empty staged word (per-quad word-occupancy mask + ffs)

  python scripts/blend_walk_sim.py [--shape replica] [--tiles 400]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from oracle import oracle as orc  # noqa: E402


def hit_masks(rec, bx0, by0):
    """[n, 32] bool: publish_hits of blend_lq.cu for n entries against the 32 quads of the half tile at (bx0, by0)."""
    x, y, ha, b, hc, L = rec[:, 0], rec[:, 1], rec[:, 2], rec[:, 3], rec[:, 4], rec[:, 5]
    a, c = 2.0 * ha, 2.0 * hc
    inv_det = 1.0 / (a * c - b * b)
    twoL = 2.0 * L * inv_det
    cxx, cyy, cxy = c * twoL, a * twoL, np.abs(b * twoL)
    hx2 = cxx * 1.01 + 0.05
    hy2 = cyy * 1.01 + 0.05
    R2 = (np.maximum(cxx, cyy) + cxy) * 1.01 + 0.05
    X, Y = x - bx0, y - by0
    cols = np.arange(8, dtype=np.float32)
    rows = np.arange(4, dtype=np.float32)
    ex = np.maximum(np.maximum(2 * cols[None] - X[:, None], X[:, None] - (2 * cols[None] + 1)), 0.0)
    ey = np.maximum(np.maximum(2 * rows[None] - Y[:, None], Y[:, None] - (2 * rows[None] + 1)), 0.0)
    ex2, ey2 = ex * ex, ey * ey
    thr = np.where(ey2 <= hy2[:, None], np.minimum(hx2[:, None], R2[:, None] - ey2), -1.0)   # [n,4]
    hit = ex2[:, None, :] <= thr[:, :, None]                                                  # [n,4,8]
    return hit.reshape(-1, 32)


def walk(hits_per_lane, n_entries, nw, skip):
    """hits_per_lane: list of 32 sorted int arrays (entry positions in walk order).  Returns (iterations, live)."""
    n_words = (n_entries + 31) >> 5
    if n_words == 0:
        return 0, 0
    ptr = np.zeros(32, np.int64)
    nh = np.array([len(h) for h in hits_per_lane])
    # next-hit position per lane (inf when exhausted)
    pad = [np.concatenate([h, [1 << 40]]) for h in hits_per_lane]
    done_lane = nh == 0
    staged, base = min(nw, n_words), 0
    # the kernel starts a lane at the word of its first candidate entry (e_first >> 5), capped by the staged window
    jq = np.minimum(np.array([p[0] for p in pad]) >> 5, staged - 1)
    iters = live = 0
    FAR = 1 << 30
    while True:
        nxt = np.array([pad[l][ptr[l]] for l in range(32)])
        nxt_word = nxt >> 5
        in_ring = nxt < staged * 32
        if skip:
            # jump to the word of the next staged hit, or to the last staged word when none is staged
            jq = np.where(in_ring, nxt_word, np.maximum(jq, staged - 1))
        else:
            # one empty word per iteration: advance only when the current word holds no further hit of this lane
            cur_has = in_ring & (nxt_word == jq)
            adv = (~cur_has) & (jq + 1 < staged)
            jq = jq + adv
            cur_has = in_ring & (nxt_word == jq)
            in_ring = cur_has
        act = in_ring & (nxt_word == jq) & ~done_lane
        eff = np.where(done_lane, FAR, np.where(act, jq, jq + 1))
        wmin = eff.min()
        if wmin >= n_words:
            break
        iters += 1
        live += int(act.sum())
        ptr = ptr + act
        done_lane = done_lane | (ptr >= nh)
        if wmin > base:
            base += 1
            if staged < n_words:
                staged += 1
    return iters, live


def walk_shared(hits_a, hits_b, n_entries, nw):
    """Two warps (lists of 32 hit arrays each) over ONE ring of nw words: a word is recycled when both warps are past
    it.  The warps are advanced in lock step (one iteration each per round).  Returns (iterations a+b, live a+b)."""
    n_words = (n_entries + 31) >> 5
    if n_words == 0:
        return 0, 0
    W = []
    for hits in (hits_a, hits_b):
        nh = np.array([len(h) for h in hits])
        pad = [np.concatenate([h, [1 << 40]]) for h in hits]
        W.append(dict(ptr=np.zeros(32, np.int64), nh=nh, pad=pad, done=nh == 0, base=0, fin=bool((nh == 0).all()),
                      jq=np.minimum(np.array([p[0] for p in pad]) >> 5, min(nw, n_words) - 1)))
    staged = min(nw, n_words)
    iters = live = 0
    FAR = 1 << 30
    guard = 0
    while not (W[0]["fin"] and W[1]["fin"]):
        guard += 1
        if guard > 100000:
            break
        for w in W:
            if w["fin"]:
                continue
            nxt = np.array([w["pad"][l][w["ptr"][l]] for l in range(32)])
            nxt_word = nxt >> 5
            in_ring = nxt < staged * 32
            cur_has = in_ring & (nxt_word == w["jq"])
            adv = (~cur_has) & (w["jq"] + 1 < staged)
            w["jq"] = w["jq"] + adv
            act = in_ring & (nxt_word == w["jq"]) & ~w["done"]
            eff = np.where(w["done"], FAR, np.where(act, w["jq"], w["jq"] + 1))
            wmin = eff.min()
            if wmin >= n_words:
                w["fin"] = True
                w["base"] = FAR
                continue
            iters += 1
            live += int(act.sum())
            w["ptr"] = w["ptr"] + act
            w["done"] = w["done"] | (w["ptr"] >= w["nh"])
            if wmin > w["base"]:
                w["base"] += 1
        base = min(W[0]["base"], W[1]["base"])
        staged = min(n_words, max(staged, min(base + nw, n_words)))
    return iters, live


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="replica")
    ap.add_argument("--tiles", type=int, default=300, help="random sample of tiles to simulate")
    ap.add_argument("--n", type=int, default=None)
    a = ap.parse_args()
    n_gauss = a.n or {"tum": 307200, "scannetpp": 3000000}.get(a.shape, 1000000)
    syn, params, kfs = bench.build_workload(a.shape, n_gauss, 2)
    W, H, fx, fy, cx, cy = syn.SHAPES[a.shape][:6]
    ocam = orc.make_camera(W, H, fx, fy, cx, cy)
    k = kfs[1]
    flags = orc.F_LOG_SCALES | orc.F_LOGIT_OPAC
    o = orc.render_full(ocam, params["means3D"], params["log_scales"], None, params["logit_opacities"].reshape(-1),
                        params["rgb_colors"], ch=5, flags=flags, pose=orc.make_pose(k["q"], k["t"]))
    rec, vals, ranges, last = o["rec"], o["vals"], o["ranges"], o["last"]
    tw, th = (W + 15) // 16, (H + 15) // 16
    rng = np.random.RandomState(0)
    tiles = rng.choice(tw * th, min(a.tiles, tw * th), replace=False)
    variants = [("nw8", 8, False, False), ("nw16", 16, False, False), ("nw32", 32, False, False), ("nw64", 64, False, False), ("nw64skip", 64, False, True)]
    tot = {v[0]: [0, 0, 0] for v in variants}   # iterations, live, staged words
    pairs = 0
    for t in tiles:
        ty, tx = divmod(int(t), tw)
        s, e = int(ranges[t]), int(ranges[t + 1])
        if e <= s:
            continue
        # tile-shared ring: both halves walk from the TILE's last contributor
        blk = last[ty * 16:ty * 16 + 16, tx * 16:tx * 16 + 16]
        tile_last = int(blk.max()) if blk.size else s - 1
        if tile_last >= s:
            n_t = tile_last - s + 1
            ids_t = vals[s:tile_last + 1][::-1]
            hh = []
            for half in (0, 1):
                x0, y0 = tx * 16, ty * 16 + 8 * half
                b2 = last[y0:y0 + 8, x0:x0 + 16]
                full = np.full((8, 16), s - 1, np.int64)
                full[:b2.shape[0], :b2.shape[1]] = b2
                ql = full.reshape(4, 2, 8, 2).max(axis=(1, 3)).reshape(32)
                mt = hit_masks(rec[ids_t], x0 + 0.5, y0 + 0.5)
                mt = mt & (np.arange(n_t)[:, None] >= (tile_last - ql)[None, :]) & (ql[None, :] >= s)
                hh.append([np.nonzero(mt[:, q])[0] for q in range(32)])
            for nw in (16, 24):
                it, lv = walk_shared(hh[0], hh[1], n_t, nw)
                key = f"tile-shared nw{nw}"
                if key not in tot:
                    tot[key] = [0, 0, 0]
                tot[key][0] += it
                tot[key][1] += lv
                tot[key][2] += (n_t + 31) >> 5
        for half in (0, 1):
            x0, y0 = tx * 16, ty * 16 + 8 * half
            blk = last[y0:y0 + 8, x0:x0 + 16]
            if blk.size == 0:
                continue
            full = np.full((8, 16), s - 1, np.int64)
            full[:blk.shape[0], :blk.shape[1]] = blk
            qlast = full.reshape(4, 2, 8, 2).max(axis=(1, 3)).reshape(32)          # last contributor per quad
            warp_last = int(qlast.max())
            if warp_last < s:
                continue
            n_entries = warp_last - s + 1
            ids = vals[s:warp_last + 1][::-1]                                        # backward walk order
            m = hit_masks(rec[ids], x0 + 0.5, y0 + 0.5)
            e_first = warp_last - qlast                                             # entries in front are skipped
            pos = np.arange(n_entries)
            m = m & (pos[:, None] >= e_first[None, :]) & (qlast[None, :] >= s)
            pairs += int(m.sum())
            # chunked variant: the CTA stages C source words, every quad walks its hits of the chunk, barrier, next chunk
            for C in (8, 16, 24, 32):
                nchunk = (n_entries + 32 * C - 1) // (32 * C)
                it = lv = 0
                for c in range(nchunk):
                    cnt = m[c * 32 * C:(c + 1) * 32 * C].sum(0)
                    it += int(cnt.max())
                    lv += int(cnt.sum())
                key = f"chunk{C}"
                if key not in tot:
                    tot[key] = [0, 0, 0]
                tot[key][0] += it
                tot[key][1] += lv
                tot[key][2] += (n_entries + 31) >> 5
            for name, nw, compact, skip in variants:
                if compact:
                    keep = m.any(1)
                    # the ring must still start at entry 0 of the walk: compaction only drops rows
                    mm = m[keep]
                else:
                    mm = m
                hits = [np.nonzero(mm[:, q])[0] for q in range(32)]
                it, lv = walk(hits, mm.shape[0], nw, skip)
                tot[name][0] += it
                tot[name][1] += lv
                tot[name][2] += (mm.shape[0] + 31) >> 5
    scale = (tw * th) / len(tiles)
    print(f"shape {a.shape}: {len(tiles)} tiles sampled of {tw * th}; (quad, Gaussian) pairs extrapolated {pairs * scale / 1e6:.2f} M")
    for name in tot:
        it, lv, wd = tot[name]
        print(f"{name:22s} warp iterations {it * scale / 1e6:6.3f} M   live lanes / iteration {lv / max(it, 1):5.2f}   "
              f"staged words {wd * scale / 1e3:7.1f} K")


if __name__ == "__main__":
    main()
