"""Host plumbing for more than one rank: one process per GPU, rendezvous through torch.distributed.

The ghost exchange itself runs inside the CUDA library (NCCL send/recv on the library's stream, see
csrc/runtime_comm.inc); torch.distributed only broadcasts the NCCL unique id.  With the gloo backend
(CPU tests of the host logic against the test-only emulation build) the exchange and the all-reduce go
through callbacks into torch.distributed instead.
"""
import ctypes as C
import os
import sys

import numpy as np

_EXCHANGE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.POINTER(C.c_double)),
                           C.POINTER(C.c_longlong), C.POINTER(C.POINTER(C.c_double)), C.POINTER(C.c_longlong), C.c_void_p)
_ALLREDUCE_FN = C.CFUNCTYPE(None, C.POINTER(C.c_double), C.c_int, C.c_int, C.c_void_p)


def init_nccl(lib, p2p=None, segment_bytes=32 << 20):
    """Creates the library's NCCL communicator over the ranks of the default torch.distributed group and, unless
    p2p is False (or HB_P2P=0), the peer-memory windows for the ghost exchange (all ranks on one node)."""
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    ident = (C.c_char * 128)()
    if rank == 0:
        lib.check(lib.dll.hb_nccl_unique_id(C.cast(ident, C.c_void_p)))
    t = torch.tensor(list(ident.raw), dtype=torch.uint8, device="cuda")
    dist.broadcast(t, 0)
    raw = bytes(t.cpu().tolist())
    buf = (C.c_char * 128).from_buffer_copy(raw)
    lib.check(lib.dll.hb_nccl_init(C.cast(buf, C.c_void_p), rank, world))
    if p2p is None:
        p2p = os.environ.get("HB_P2P", "1") != "0"
    if p2p and world > 1:
        init_p2p(lib, segment_bytes)


def init_p2p(lib, segment_bytes=32 << 20):
    """Peer-memory windows: all-gather of the CUDA IPC handles, then every rank maps its neighbours' windows.
    Falls back (on every rank together, with a message) to the NCCL send/recv exchange if any rank cannot map."""
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    handle = (C.c_char * 64)()
    ok = lib.dll.hb_p2p_window(C.cast(handle, C.c_void_p), segment_bytes, world) == 0
    mine = torch.tensor(list(handle.raw), dtype=torch.uint8, device="cuda")
    every = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(every, mine)
    if ok:
        raw = b"".join(bytes(t.cpu().tolist()) for t in every)
        buf = (C.c_char * (64 * world)).from_buffer_copy(raw)
        ok = lib.dll.hb_p2p_connect(C.cast(buf, C.c_void_p), rank, world) == 0
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if int(flag.item()) == 0:
        lib.dll.hb_p2p_finalize()
        if rank == 0:
            print("hyteg_b200: peer-memory windows unavailable (%s); ghost exchange stays on NCCL send/recv" % lib.dll.hb_last_error().decode(), file=sys.stderr)
        return False
    return True


def p2p_active(lib):
    n = C.c_longlong(0)
    return bool(lib.dll.hb_p2p_active(C.byref(n))), n.value


def init_callbacks(lib):
    """Exchange + all-reduce through torch.distributed point-to-point calls on HOST buffers (gloo, emulation build)."""
    import torch
    import torch.distributed as dist

    def exchange(ctx, n_peers, peers, send, n_send, recv, n_recv, stream):
        try:
            works, keep = [], []
            for k in range(n_peers):
                if n_recv[k] > 0:
                    b = torch.from_numpy(np.ctypeslib.as_array(recv[k], shape=(n_recv[k],)))
                    keep.append(b)
                    works.append(dist.irecv(b, peers[k]))
            for k in range(n_peers):
                if n_send[k] > 0:
                    a = torch.from_numpy(np.ctypeslib.as_array(send[k], shape=(n_send[k],)))
                    keep.append(a)
                    works.append(dist.isend(a, peers[k]))
            for w in works:
                w.wait()
            return 0
        except Exception as e:  # pragma: no cover
            print("hyteg_b200 exchange callback failed:", e)
            return 1

    def allreduce(buf, n, op, ctx):
        t = torch.from_numpy(np.ctypeslib.as_array(buf, shape=(n,)))
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == 2 else dist.ReduceOp.SUM)

    ex, ar = _EXCHANGE_FN(exchange), _ALLREDUCE_FN(allreduce)
    lib._keepalive += [ex, ar]
    lib.check(lib.dll.hb_set_exchange(C.cast(ex, C.c_void_p), None))
    lib.check(lib.dll.hb_set_allreduce(C.cast(ar, C.c_void_p), None))


def exchange_stats(lib):
    a, b = C.c_longlong(0), C.c_longlong(0)
    lib.dll.hb_exchange_stats(C.byref(a), C.byref(b))
    return a.value, b.value
