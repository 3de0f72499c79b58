import sys, time, numpy as np, torch
sys.path.insert(0, ".")
import bench_config as cfg
sc = cfg.two_view_scene()
p1 = np.ascontiguousarray(sc["x1"], np.float64); p2 = np.ascontiguousarray(sc["x2"], np.float64)
n = len(p1)
pin = torch.empty((2 * n, 2), dtype=torch.float64, pin_memory=True); hx = pin.numpy()
d = torch.empty((2 * n, 2), dtype=torch.float64, device="cuda")
def t(f, name, reps=20):
    f(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a = time.perf_counter(); f(); ts.append(time.perf_counter() - a)
    torch.cuda.synchronize()
    print(f"{name}: host-side min {1e3*min(ts):.3f} ms median {1e3*sorted(ts)[reps//2]:.3f} ms")
def v1():
    np.copyto(hx[:n], p1); np.copyto(hx[n:], p2); d.copy_(pin, non_blocking=True)
def v2():
    d[:n].copy_(torch.from_numpy(p1), non_blocking=True); d[n:].copy_(torch.from_numpy(p2), non_blocking=True)
def v3():
    hx[:n] = p1; hx[n:] = p2; d.copy_(pin, non_blocking=True)
import ctypes
def v4():
    ctypes.memmove(hx.ctypes.data, p1.ctypes.data, p1.nbytes); ctypes.memmove(hx.ctypes.data + p1.nbytes, p2.ctypes.data, p2.nbytes); d.copy_(pin, non_blocking=True)
t(v1, "copyto pinned + async H2D")
t(v2, "pageable H2D x2")
t(v3, "slice assign pinned + async H2D")
t(v4, "memmove pinned + async H2D")
plain = np.empty((2 * n, 2))
def v5():
    np.copyto(plain[:n], p1); np.copyto(plain[n:], p2)
t(v5, "copyto pageable numpy (no CUDA)")
