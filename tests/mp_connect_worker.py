"""Worker for tests/test_gpu_exchange.py::test_ipc_rendezvous_ignores_strangers (torchrun, 2 ranks,
both on cuda:0: only the rendezvous runs, no kernel that waits for a peer).

Before connecting, every rank fires datagrams at its OWN rendezvous socket the way a stranger on
the host would: right magic, a plausible rank, a file descriptor attached -- but without the
64-bit secret that only travels inside the exported handles.  Round 1 accepted such a message and
would have mapped the stranger's file as the peer's segment; now it must be dropped and the real
peers must still connect."""
import os
import socket
import struct
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "kind-gpu-sim_b200")):
    sys.path.insert(0, p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import gpuprobe  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]) if os.environ.get("GPU_PROBE_CONNECT_TWO_GPUS") else 0)
    dist.init_process_group("gloo")
    x = gpuprobe.Exchange(rank, world, 1 << 20)
    mine = x.export_handle()
    magic, hrank, mc, pid, seg_bytes, token = struct.unpack_from("<IiiiQQ", mine, 0)
    name = mine[32:64].split(b"\0", 1)[0]
    assert hrank == rank and pid == os.getpid() and token != 0 and name.startswith(b"gpuprobe-")
    handles = [None] * world
    dist.all_gather_object(handles, mine)
    assert len({struct.unpack_from("<Q", h, 24)[0] for h in handles}) == world  # every rank its own secret

    # the stranger: MSG_MEM (type 1) "from" the other rank, wrong secret, a real fd (/dev/null) attached
    rogue = socket.socket(socket.AF_UNIX, socket.SOCK_DGRAM)
    other = (rank + 1) % world
    devnull = os.open("/dev/null", os.O_RDONLY)
    for bad_token in (0, token ^ 1, struct.unpack_from("<Q", handles[other], 24)[0]):  # none, nearly, the PEER's secret
        msg = struct.pack("<IiiiQ", 0x47505845, 1, other, 0, bad_token)
        rogue.sendmsg([msg], [(socket.SOL_SOCKET, socket.SCM_RIGHTS, struct.pack("i", devnull))], 0, b"\0" + name)
    rogue.sendmsg([b"garbage"], [], 0, b"\0" + name)  # and something that is not a message at all
    os.close(devnull)

    x.connect_ipc(b"".join(handles))
    dist.barrier()
    ok = x.status() == 0
    x.close()
    dist.barrier()
    if rank == 0 and ok:
        print(f"CONNECT_OK world={world}", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
