"""Do a pinned H2D copy and a pinned D2H copy on two streams overlap on this box?  Times (CUDA events) the frame upload of a
250-frame batch alone, the result download alone and both at once; under torchrun every rank does it at the same time.
usage: python tools/copy_probe.py   |   torchrun --nproc-per-node N tools/copy_probe.py"""
import json
import os

import torch
import torch.distributed as dist


def main():
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    up_b, dn_b = 250 * 752 * 480, 250 * 1000 * 65
    hu = torch.empty(up_b, dtype=torch.uint8).pin_memory(); du = torch.empty(up_b, dtype=torch.uint8, device="cuda")
    hd = torch.empty(dn_b, dtype=torch.uint8).pin_memory(); dd = torch.empty(dn_b, dtype=torch.uint8, device="cuda")
    su, sd = torch.cuda.Stream(), torch.cuda.Stream()

    def run(up, dn, reps=8):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        eu = torch.cuda.Event(enable_timing=True); ed = torch.cuda.Event(enable_timing=True)
        e0.record()
        su.wait_event(e0); sd.wait_event(e0)
        for _ in range(reps):
            if up:
                with torch.cuda.stream(su):
                    du.copy_(hu, non_blocking=True)
            if dn:
                with torch.cuda.stream(sd):
                    hd.copy_(dd, non_blocking=True)
        eu.record(su); ed.record(sd)
        torch.cuda.current_stream().wait_event(eu); torch.cuda.current_stream().wait_event(ed)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(eu) / reps, e0.elapsed_time(ed) / reps, e0.elapsed_time(e1) / reps

    run(True, True, 2)
    a = run(True, False); b = run(False, True); c = run(True, True)
    res = {"rank": rank, "world": world, "h2d_alone_ms": round(a[0], 3), "h2d_alone_GBps": round(up_b / a[0] / 1e6, 2),
           "d2h_alone_ms": round(b[1], 3), "d2h_alone_GBps": round(dn_b / b[1] / 1e6, 2),
           "both_h2d_ms": round(c[0], 3), "both_d2h_ms": round(c[1], 3), "both_total_ms": round(c[2], 3)}
    if world > 1:
        allr = [None] * world
        dist.all_gather_object(allr, res)
        if rank == 0:
            for r in allr:
                print(json.dumps(r))
        dist.destroy_process_group()
    else:
        print(json.dumps(res))


if __name__ == "__main__":
    main()
