"""Transport of the multi-GPU exchanges.  What a rank contributes to an exchange is packed ON THE
DEVICE into one byte message by the CUDA side (csrc/exchange.cu, csrc/stab.cu); this module all-gathers such messages
— torch.distributed over NCCL/NVLink on the GPU box (the bytes never visit the host), gloo in the CPU tests, plain
hand-over between thread ranks — and hands the gathered buffer back to the CUDA side.

PyTorch is plumbing here: device memory views and torch.distributed.  The arithmetic of the path stays in the CUDA
kernels behind the C ABI.
"""
import numpy as np
import torch

EAGER = 256 * 1024     # bytes per rank of the first (fixed-size) round; larger messages take a second, exact round
_HDR = 16              # per-rank prefix of the first round: int64 message size, 8 bytes of padding


class DevPtr(object):
    """A raw device address as a __cuda_array_interface__ object, so torch can view it without a copy."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def view(ptr, n, typestr, device):
    if n == 0:
        return torch.zeros(0, dtype={"|u1": torch.uint8, "<i8": torch.int64, "<i4": torch.int32}[typestr], device=device)
    return torch.as_tensor(DevPtr(ptr, (n,), typestr), device=device)


def _pad16(n):
    return (n + 15) & ~15


class TensorComm(object):
    """All-gather of one variable-length byte message per rank, and integer all-reduce, over a torch.distributed
    group, in-process thread ranks (sharded.ThreadComm) or a single rank.

    gather_msgs costs ONE collective and one host read of the sizes when every message fits the eager slot (the
    usual case: a few thousand rows), two collectives otherwise.  No size exchange precedes it: the first round has a
    fixed shape, each rank's slot carrying its message size and the head of the message."""

    def __init__(self, comm, eager=EAGER):
        self._c = comm
        self.rank, self.world = comm.rank, comm.world
        self._dist = getattr(comm, "_dist", None)
        self._group = getattr(comm, "_group", None)
        self._eager = int(eager)
        self._send = None
        self.n_collectives = 0

    def _buffers(self, device):
        if self._send is None or self._send.device != device:
            self._send = torch.zeros(self._eager, dtype=torch.uint8, device=device)
            self._size = torch.zeros(1, dtype=torch.int64, device=device)
        # the gathered buffer is handed to the caller, who may still read it after the next gather: a fresh one per call
        return self._send, torch.empty(self.world * self._eager, dtype=torch.uint8, device=device)

    def gather_msgs(self, msg):
        """msg: uint8 tensor [nbytes], nbytes % 16 == 0.  -> (buf uint8 tensor, starts, sizes): message r of the
        group is buf[starts[r] : starts[r] + sizes[r]]; every start is 16-byte aligned; same on every rank."""
        n = int(msg.shape[0])
        assert n % 16 == 0, "messages are padded to 16 bytes"
        if self.world == 1:
            return msg, [0], [n]
        if self._dist is None:                      # thread ranks share the process and the device: hand the tensors over
            parts = self._c.all_gather(msg)
            sizes = [int(p.shape[0]) for p in parts]
            starts = np.concatenate([[0], np.cumsum(sizes)[:-1]]).tolist()
            buf = torch.cat(parts)
            if buf.is_cuda:
                torch.cuda.current_stream().synchronize()
            return buf, [int(s) for s in starts], sizes
        send, recv = self._buffers(msg.device)
        room = self._eager - _HDR
        self._size[0] = n
        send[:8] = self._size.view(torch.uint8)
        if n <= room:
            send[_HDR:_HDR + n] = msg
        self._dist.all_gather_into_tensor(recv, send, group=self._group)
        self.n_collectives += 1
        sizes = recv.view(self.world, self._eager)[:, :8].contiguous().view(torch.int64).flatten().tolist()   # host sync
        if max(sizes) <= room:
            return recv, [r * self._eager + _HDR for r in range(self.world)], sizes
        stride = _pad16(max(sizes))
        pad = torch.empty(stride, dtype=torch.uint8, device=msg.device)
        pad[:n] = msg
        out = torch.empty(self.world * stride, dtype=torch.uint8, device=msg.device)
        self._dist.all_gather_into_tensor(out, pad, group=self._group)
        self.n_collectives += 1
        if out.is_cuda:
            torch.cuda.current_stream().synchronize()
        return out, [r * stride for r in range(self.world)], sizes

    def all_reduce_sum(self, t):
        if self.world == 1:
            return t
        if self._dist is None:
            return torch.as_tensor(self._c.all_reduce_sum(t.cpu().numpy())).to(t.device)
        t = t.clone()
        self._dist.all_reduce(t, op=self._dist.ReduceOp.SUM, group=self._group)
        self.n_collectives += 1
        return t


def host_msg(arr, device):
    """A host-built message (numpy array of any dtype) as a 16-byte padded uint8 tensor on `device`."""
    raw = np.ascontiguousarray(arr).view(np.uint8).reshape(-1)
    out = np.zeros(_pad16(len(raw)), np.uint8)
    out[:len(raw)] = raw
    return torch.from_numpy(out).to(device)


def row_names(text, off):
    """Read names of the rows of a rows-only ctx: every text slot starts with the NUL-terminated name."""
    text = np.asarray(text, dtype=np.uint8)
    off = np.asarray(off, dtype=np.int64)
    if len(off) == 0:
        return []
    nul = np.flatnonzero(text == 0)
    end = nul[np.searchsorted(nul, off)]
    raw = text.tobytes()
    return [raw[a:b].decode() for a, b in zip(off.tolist(), end.tolist())]
