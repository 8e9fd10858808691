"""Debug aid: dump the matcher's intermediate volumes for one small pair and compare them with numpy."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import stereo.vision_b200._native as N
import stereo.vision_b200.synth as synth
import sgbm_oracle as so

H, W, D = 24, 300, 128
L, R = synth.make_stereo_pair(2, 5, (120, 400))
L, R = np.ascontiguousarray(L[:H, :W]), np.ascontiguousarray(R[:H, :W])
lib = N.load()
cuda = torch.device("cuda:0")
nb = lib.svb_sgbm_scratch_bytes(1, H, W, 128)
scratch = torch.zeros((nb,), dtype=torch.uint8, device=cuda)
out = torch.zeros((1, H, W), dtype=torch.int16, device=cuda)
dL, dR = torch.from_numpy(L).to(cuda), torch.from_numpy(R).to(cuda)
N.check(lib.svb_sgbm_compute(N.ptr(dL), N.ptr(dR), 1, H, W, 128, 21, N.ptr(out), N.ptr(scratch), nb, 0), "compute")
torch.cuda.synchronize()
off = (-scratch.data_ptr()) % 256
s = scratch[off:].cpu().numpy()
n = H * W; w1 = W - D
planes = s[:6 * n * 4].view(np.uint32).reshape(2, 3, H, W)
vol = H * w1 * D * 2
S = s[6 * n * 4:6 * n * 4 + vol].view(np.int16).reshape(H, w1, D)
C = s[6 * n * 4 + vol:6 * n * 4 + 2 * vol].view(np.int16).reshape(H, w1, D)
raw = s[6 * n * 4 + 2 * vol:6 * n * 4 + 2 * vol + n * 2].view(np.int16).reshape(H, W)
print("planes L value g/r row 3:", planes[0, 0, 3, 128:134] & 0xffff, planes[0, 0, 3, 128:134] >> 16, "raw L", L[3, 128:134])
print("C stats", C.min(), C.max(), C.mean(), "S stats", S.min(), S.max(), "raw", np.unique(raw)[:10])
# numpy pixel cost
def prefilter(img):
    I = img.astype(np.int32); Hh, Ww = I.shape
    up = np.vstack([I[:1], I[:-1]]); dn = np.vstack([I[1:], I[-1:]])
    g = np.full_like(I, 15); r = np.full_like(I, 15)
    gx = lambda A: A[:, 2:] - A[:, :-2]
    g[:, 1:-1] = np.clip(gx(I) * 2 + gx(up) + gx(dn), -15, 15) + 15
    r[:, 1:-1] = I[:, 1:-1]
    return g, r
def interval(p):
    l = np.concatenate([p[:, :1], (p[:, 1:] + p[:, :-1]) // 2], 1); r_ = np.concatenate([(p[:, :-1] + p[:, 1:]) // 2, p[:, -1:]], 1)
    return np.minimum(p, np.minimum(l, r_)), np.maximum(p, np.maximum(l, r_))
pix = np.zeros((H, w1, D), np.int32)
for (pl, pr, sh) in [(prefilter(L)[0], prefilter(R)[0], 0), (prefilter(L)[1], prefilter(R)[1], 2)]:
    l0, l1 = interval(pl); r0, r1 = interval(pr)
    for d in range(D):
        u, u0, u1 = pl[:, D:], l0[:, D:], l1[:, D:]
        v, v0, v1 = pr[:, D - d:W - d], r0[:, D - d:W - d], r1[:, D - d:W - d]
        c0 = np.maximum(0, np.maximum(u - v1, v0 - u)); c1 = np.maximum(0, np.maximum(v - u1, u0 - v))
        pix[:, :, d] += np.minimum(c0, c1) >> sh
idx = np.clip(np.arange(-10, w1 + 10), 0, w1 - 1)
hs = sum(pix[:, idx[k:k + w1], :] for k in range(21))
idy = np.clip(np.arange(-10, H + 10), 0, H - 1)
Cn = sum(hs[idy[k:k + H]] for k in range(21)).astype(np.int16)
print("C equal:", np.array_equal(C, Cn), "diff count", int((C != Cn).sum()))
want_pre = so.sgbm_compute(L, R, 128, 21, median=False)
print("raw (pre-median) equal:", np.array_equal(raw, want_pre), int((raw != want_pre).sum()))
print("final equal:", np.array_equal(out[0].cpu().numpy(), so.sgbm_compute(L, R)))
ys, xs = np.nonzero(raw != want_pre)
for y, x in list(zip(ys, xs))[:8]: print(y, x, raw[y, x], want_pre[y, x])
