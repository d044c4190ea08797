import sys, torch
sys.path.insert(0, '/root/repo')
from woft_b200 import least_squares_H as lsq
dev = 'cuda:0'
g = torch.Generator(device=dev).manual_seed(0)
Bh, Nh = 1024, 14400
p1 = torch.rand((Bh, Nh, 2), generator=g, device=dev) * torch.tensor([1280.0, 720.0], device=dev)
p2 = p1 + torch.randn((Bh, Nh, 2), generator=g, device=dev)
w = torch.rand((Bh, Nh), generator=g, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for name, fn in (("qr", lambda: lsq.find_homography_nonhomogeneous_QR(p1, p2, w)), ("dlt", lambda: lsq.find_homography_dlt(p1, p2, w)),
                 ("irls_l1", lambda: lsq.find_homography_IRLSq_QR(p1, p2, w, reweighting_fn=lsq.IRLSq_L1, n_iter=5))):
    for _ in range(2): fn()
    ts = []
    for _ in range(5):
        flush.zero_(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); b.synchronize(); ts.append(a.elapsed_time(b) * 1e3)
    print(name, f"{sorted(ts)[2]:.1f} us")
