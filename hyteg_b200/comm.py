"""Host plumbing for more than one rank: one process per GPU, rendezvous through torch.distributed.

The ghost exchange itself runs inside the CUDA library (NCCL send/recv on the library's stream, see
csrc/runtime_comm.inc); torch.distributed only broadcasts the NCCL unique id.  With the gloo backend
(CPU tests of the host logic against the test-only emulation build) the exchange and the all-reduce go
through callbacks into torch.distributed instead.
"""
import ctypes as C

import numpy as np

_EXCHANGE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.POINTER(C.c_double)),
                           C.POINTER(C.c_longlong), C.POINTER(C.POINTER(C.c_double)), C.POINTER(C.c_longlong), C.c_void_p)
_ALLREDUCE_FN = C.CFUNCTYPE(None, C.POINTER(C.c_double), C.c_int, C.c_int, C.c_void_p)


def init_nccl(lib):
    """Creates the library's NCCL communicator over the ranks of the default torch.distributed group."""
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
