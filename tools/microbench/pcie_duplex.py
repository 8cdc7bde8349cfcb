"""What the PCIe link of the box gives for the end-to-end step's copy sizes: H2D alone, D2H alone, both at once (two streams),
as one copy each and split the way DexStepEnv.step_host splits them.  Prints GB/s per direction."""
import sys
import torch

dev = "cuda:0"
H2D = [16384 * b for b in (208, 336, 72, 52, 72)]   # dof_state, rigid (packed), contact, root, actions
D2H = [16384 * b for b in (632, 4, 1)]              # obs, rew, reset


def bufs(sizes):
    return [(torch.empty(s, dtype=torch.uint8, pin_memory=True), torch.empty(s, dtype=torch.uint8, device=dev)) for s in sizes]


def run(h_sizes, d_sizes, iters=200):
    hb, db = bufs(h_sizes), bufs(d_sizes)
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    main = torch.cuda.current_stream(dev)

    def go(k):
        for _ in range(k):
            with torch.cuda.stream(s1):
                for h, d in hb:
                    d.copy_(h, non_blocking=True)
            with torch.cuda.stream(s2):
                for h, d in db:
                    h.copy_(d, non_blocking=True)

    go(10)
    torch.cuda.synchronize()
    e0.record(main)
    s1.wait_event(e0)
    s2.wait_event(e0)
    go(iters)
    main.wait_stream(s1)
    main.wait_stream(s2)
    e1.record(main)
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    return us, sum(h_sizes) / us / 1e3, sum(d_sizes) / us / 1e3


for name, h, d in (("H2D alone, split", H2D, []), ("H2D alone, one copy", [sum(H2D)], []), ("D2H alone, split", [], D2H),
                   ("both, split", H2D, D2H), ("both, one copy each", [sum(H2D)], [sum(D2H)])):
    us, gh, gd = run(h, d)
    print(f"{name:24s} {us:8.1f} us per round   H2D {gh:5.1f} GB/s   D2H {gd:5.1f} GB/s")
    sys.stdout.flush()
