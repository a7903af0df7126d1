"""Worker for tests/test_gpu_exchange.py::test_exchange_across_processes_over_ipc.

Launched by torchrun, one process per GPU.  The ranks trade the 64-byte handles of their
mailboxes over gloo (host plumbing only), then run the fused checksum + lanes exchange
and the fused triad + checksum + exchange; every rank checks the result against the
CPU oracle of the WHOLE buffer."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "kind-gpu-sim_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import gpuprobe  # noqa: E402
import oracle  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    x = gpuprobe.Exchange(rank, world, 8 << 20)
    handles = [None] * world
    dist.all_gather_object(handles, x.export_handle())
    x.connect_ipc(b"".join(handles))
    dist.barrier()  # every rank has mapped every mailbox before anyone stores into one

    n_total, first = (1 << 22) + 101, 1 << 35
    lo, hi = n_total * rank // world, n_total * (rank + 1) // world
    m = hi - lo
    st = torch.cuda.current_stream().cuda_stream
    b = torch.empty(m, dtype=torch.float64, device="cuda")
    c = torch.empty(m, dtype=torch.float64, device="cuda")
    a = torch.empty(m, dtype=torch.float64, device="cuda")
    ws = torch.zeros(gpuprobe.workspace_bytes() // 8, dtype=torch.int64, device="cuda")
    out = torch.zeros(2, dtype=torch.int64, device="cuda")
    gpuprobe.init_fill(b.data_ptr(), m, first + lo, gpuprobe.SEED_B, st)
    gpuprobe.init_fill(c.data_ptr(), m, first + lo, gpuprobe.SEED_C, st)
    want = oracle.stream_lanes(n_total, first)
    for k in range(6):
        out.fill_(-1)
        if k % 2 == 0:
            gpuprobe.triad(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, m, gpuprobe.VARIANT_AUTO, st)
            x.checksum_u64(a.data_ptr(), m, first + lo, out.data_ptr(), ws.data_ptr(), gpuprobe.VARIANT_AUTO, st)
        else:
            x.triad_checksum(a.data_ptr(), b.data_ptr(), c.data_ptr(), gpuprobe.SCALAR, m, first + lo, out.data_ptr(),
                             ws.data_ptr(), gpuprobe.VARIANT_AUTO, st)
        torch.cuda.synchronize()
        got = tuple(int(v) for v in out.cpu().numpy().view(np.uint64))
        if got != want or x.status() != 0:
            print(f"rank {rank} call {k}: lanes {got} != oracle {want} (status {x.status()})", flush=True)
            sys.exit(1)
    # K4 over the same segments: integer-valued operands, so every algorithm must hit the closed form exactly
    cnt = (1 << 20) + 4
    algos = [gpuprobe.ALLREDUCE_PULL_PUSH, gpuprobe.ALLREDUCE_AUTO]
    if x.has_multicast:
        algos.append(gpuprobe.ALLREDUCE_MULTIMEM)
    pat = (torch.arange(cnt, device="cuda", dtype=torch.float32) % 7 + 1)
    for algo in algos + [gpuprobe.ALLREDUCE_PUSH]:
        n_el = cnt if algo != gpuprobe.ALLREDUCE_PUSH else 1 << 18
        src = pat[:n_el] * float(rank + 1)
        gpuprobe.p2p_copy(x.buffer, src.data_ptr(), 4 * n_el, st)
        torch.cuda.synchronize()
        dist.barrier()
        for _ in range(2):
            x.allreduce_f32(0, n_el, 0, st, algo)
            outp = torch.empty(n_el, dtype=torch.float32, device="cuda")
            gpuprobe.p2p_copy(outp.data_ptr(), x.buffer, 4 * n_el, st)
            torch.cuda.synchronize()
            scale = float(world * (world + 1) // 2)
            if not torch.equal(outp, pat[:n_el] * scale) or x.status() != 0:
                print(f"rank {rank}: allreduce algo {algo} wrong (status {x.status()})", flush=True)
                sys.exit(1)
            gpuprobe.p2p_copy(x.buffer, src.data_ptr(), 4 * n_el, st)  # restore the inputs for the second call
            torch.cuda.synchronize()
            dist.barrier()
    mc = x.has_multicast
    dist.barrier()
    x.close()
    if rank == 0:
        print(f"EXCHANGE_IPC_OK world={world} multicast={int(mc)} lanes=0x{want[0]:016x},0x{want[1]:016x}", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
