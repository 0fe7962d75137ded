"""Host<->device copy bandwidth from pinned memory: each direction alone and both at once
(the ceiling of bench.py's e2e number, which moves 100 bytes in and 100 bytes out per record)."""
import json
import sys

import torch

gb = float(sys.argv[1]) if len(sys.argv) > 1 else 4.0
n = int(gb * 1e9)
h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True)
h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s1.wait_stream(torch.cuda.current_stream())
    s2.wait_stream(torch.cuda.current_stream())
    for _ in range(reps):
        fn()
    torch.cuda.current_stream().wait_stream(s1)
    torch.cuda.current_stream().wait_stream(s2)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def h2d():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


def both():
    h2d()
    d2h()


res = {"bytes_each_way": n}
for name, fn in (("h2d_alone", h2d), ("d2h_alone", d2h), ("both", both)):
    ms = timed(fn)
    res[name + "_ms"] = round(ms, 2)
    res[name + "_GBs_per_direction"] = round(n / ms / 1e6, 2)
print(json.dumps(res))
