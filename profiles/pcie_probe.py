"""Pinned-memory copy ceilings of the box: H2D alone, D2H alone, both at once (two streams), at the
byte ratio of the cfg-2 e2e step (1.283 MB up : 0.774 MB down per person).  Context for `e2e`."""
import json
import torch

up_b, down_b = 1283 * 2**18, 774 * 2**18  # ~336 MB / ~203 MB: 256 persons
hu = torch.empty(up_b, dtype=torch.uint8).pin_memory()
hd = torch.empty(down_b, dtype=torch.uint8).pin_memory()
du = torch.empty(up_b, dtype=torch.uint8, device="cuda")
dd = torch.empty(down_b, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=8):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    for s in (s1, s2):
        torch.cuda.current_stream().wait_stream(s)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps / 1e3


def up():
    with torch.cuda.stream(s1):
        du.copy_(hu, non_blocking=True)


def down():
    with torch.cuda.stream(s2):
        hd.copy_(dd, non_blocking=True)


def both():
    up()
    down()


for s in (s1, s2):
    s.wait_stream(torch.cuda.current_stream())
t_up, t_down, t_both = timed(up), timed(down), timed(both)
print(json.dumps({"h2d_alone_gbs": round(up_b / t_up / 1e9, 1), "d2h_alone_gbs": round(down_b / t_down / 1e9, 1),
                  "both_h2d_gbs": round(up_b / t_both / 1e9, 1), "both_d2h_gbs": round(down_b / t_both / 1e9, 1),
                  "persons_per_s_ceiling": round(256 / t_both)}))
