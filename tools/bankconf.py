"""Tiny shared-memory bank-conflict estimator: wavefronts for a warp-wide LDS/STS of `width` words per lane."""
def wavefronts(addrs, width=1):
    # addrs: word address per lane (32 lanes; None = inactive). 128-bit accesses are issued per 8-lane phase,
    # 64-bit per 16-lane phase, 32-bit per 32 lanes.
    lanes_per_phase = {1: 32, 2: 16, 4: 8}[width]
    total = 0
    for p in range(0, 32, lanes_per_phase):
        banks = {}
        for a in addrs[p:p + lanes_per_phase]:
            if a is None:
                continue
            for w in range(width):
                banks.setdefault((a + w) % 32, set()).add(a + w)
        total += max((len(v) for v in banks.values()), default=0)
    return total

if __name__ == "__main__":
    # conv1_wgrad: tid = kh + 5*(chunk + 5*c); dy1 row loads: c*cs + oy*rs (oy = chunk*5 + r); x loads (oy+kh)*28
    best = []
    for rs in range(24, 41, 4):
        for cs_pad in range(0, 33, 4):
            cs = 24 * rs + cs_pad
            tot = 0
            for warp in range(8):
                for r in range(5):
                    addrs = []
                    for lane in range(32):
                        tid = warp * 32 + lane
                        if tid >= 250: addrs.append(None); continue
                        kh, chunk, c = tid % 5, (tid // 5) % 5, tid // 25
                        oy = chunk * 5 + r
                        addrs.append(c * cs + oy * rs if oy < 24 else None)
                    tot += wavefronts(addrs, 4)
            best.append((tot, rs, cs))
    best.sort()
    print("conv1_wgrad dy1 loads (wavefronts, row stride, chan stride):", best[:5], "current:", [b for b in best if b[1] == 24 and b[2] == 576])
    # conv2_wgrad a1 loads: tid = cq*50 + ci*5 + kh ; addr = ci*cs + (oy+kh)*12
    res = []
    for cs in range(144, 180, 4):
        tot = 0
        for warp in range(8):
            addrs = []
            for lane in range(32):
                tid = warp * 32 + lane
                if tid >= 250: addrs.append(None); continue
                kh, ci = tid % 5, (tid % 50) // 5
                addrs.append(ci * cs + kh * 12)
            tot += wavefronts(addrs, 4)
        res.append((tot, cs))
    res.sort()
    print("conv2_wgrad a1 loads:", res[:4], "current:", [r for r in res if r[1] == 144])

def conv1_wgrad_x(xs=28, order="kh_chunk_c"):
    tot = 0
    for warp in range(8):
        for r in range(5):
            addrs = []
            for lane in range(32):
                tid = warp * 32 + lane
                if tid >= 250: addrs.append(None); continue
                if order == "kh_chunk_c":
                    kh, chunk, c = tid % 5, (tid // 5) % 5, tid // 25
                else:
                    chunk, kh, c = tid % 5, (tid // 5) % 5, tid // 25
                oy = chunk * 5 + r
                addrs.append((oy + kh) * xs if oy < 24 else None)
            tot += wavefronts(addrs, 4)
    return tot

if __name__ == "__main__":
    for xs in (28, 32, 36, 40, 44):
        print("conv1_wgrad x loads, row stride", xs, conv1_wgrad_x(xs), conv1_wgrad_x(xs, "chunk_kh_c"), "(ideal 160 if all distinct, fewer with broadcast)")
    # build stores (float2): e = tid + 256q -> cc = e//144, p = e%144
    for rs, cs in ((24, 576), (24, 580), (28, 672), (28, 680)):
        tot = 0
        for q in range(6):
            for warp in range(8):
                addrs = []
                for lane in range(32):
                    e = warp * 32 + lane + 256 * q
                    if e >= 1440: addrs.append(None); continue
                    cc, p = e // 144, e % 144
                    py, px = p // 12, p % 12
                    addrs.append(cc * cs + 2 * py * rs + 2 * px)
                tot += 2 * wavefronts(addrs, 2)
        print("conv1_wgrad build stores rs,cs", rs, cs, tot, "ideal", 2 * 2 * 45)
