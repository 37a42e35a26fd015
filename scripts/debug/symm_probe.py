"""Does torch's symmetric memory work here?  Peer pointers, one_shot_all_reduce eager and under graph capture, timing
against NCCL all_reduce for the step's message sizes."""
import os, sys, time
import torch, torch.distributed as dist
import torch.distributed._symmetric_memory as sm
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
os.makedirs("gpurun_out", exist_ok=True)
LOG = open(f"gpurun_out/symm_probe_rank{rank}.log", "w")
def say(*a):
    print(*a, file=LOG, flush=True)
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"])); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
g = dist.group.WORLD
say("pg up; backend", sm.get_backend(dev) if hasattr(sm, "get_backend") else None)
try:
    t = sm.empty(32768, dtype=torch.float32, device=dev)
    say("empty ok")
    h = sm.rendezvous(t, g)
    say("rendezvous ok", [hex(p) for p in h.buffer_ptrs])
    if rank == 0:
        print("rendezvous ok: world", h.world_size, "rank", h.rank, "ptrs", [hex(p) for p in h.buffer_ptrs][:3],
              "signal pad", h.signal_pad_size, "multicast", h.has_multicast_support, flush=True)
    t.fill_(rank + 1.0)
    torch.cuda.synchronize(); say("fill ok")
    h.barrier()
    torch.cuda.synchronize(); say("barrier ok")
    out = torch.empty_like(t)
    torch.ops.symm_mem.one_shot_all_reduce_out(t, "sum", g.group_name, out)
    torch.cuda.synchronize(); say("one_shot eager", float(out[0]))
    if rank == 0:
        print("one_shot eager:", float(out[0]), "expect", world * (world + 1) / 2, flush=True)
    def timeit(fn, n=200):
        for _ in range(10): fn()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(n): fn()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / n * 1e3
    for n in (1024, 8192, 30000):
        v = t[:n]; o = out[:n]; w = torch.ones(n, device=dev)
        us1 = timeit(lambda: torch.ops.symm_mem.one_shot_all_reduce_out(v, "sum", g.group_name, o))
        say("timed one_shot", n, us1)
        us2 = timeit(lambda: dist.all_reduce(w))
        say("timed nccl", n, us2)
        # graph capture of both
        res = {}
        for name, fn in (("one_shot", lambda: torch.ops.symm_mem.one_shot_all_reduce_out(v, "sum", g.group_name, o)),
                         ("nccl", lambda: dist.all_reduce(w))):
            try:
                gr = torch.cuda.CUDAGraph()
                s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(s):
                    fn()
                torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
                with torch.cuda.graph(gr):
                    for _ in range(10): fn()
                res[name] = timeit(gr.replay, 50) / 10
                say("graph", name, n, res[name])
            except Exception as e:
                res[name] = "ERR " + repr(e)[:80]
        if rank == 0:
            print(f"n={n}: eager one_shot {us1:.1f} us, nccl {us2:.1f} us; in-graph per call: {res}", flush=True)
except Exception as e:
    import traceback; say(traceback.format_exc())
dist.destroy_process_group()
