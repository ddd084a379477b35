"""Flat parameter / gradient / buffer arenas and the bucket plan.

All parameters of a model live in ONE fp32 arena and all gradients in another, laid out in
reverse registration order (roughly the order gradients become ready in backward) and cut into
buckets: a small first bucket so communication starts early, then ``bucket_cap_mb`` ones -- the
same policy DistributedDataParallel uses after its first-iteration rebuild
(the reference wraps with DDP at
pytorch/distributed_data_parallel.py:74).  ``p.data`` and the kernels' gradient targets are
views of the arenas, so "copy grads into a bucket" and "copy the result back" do not exist.

On the ``b200`` backend the arenas are symmetric peer-memory buffers, so a collective kernel on
any GPU can address every replica's copy directly.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

ALIGN = 8  # elements: every tensor 32-byte aligned in the fp32 arenas, 16-byte in a bf16 gradient arena


@dataclass
class Bucket:
    index: int
    offset: int      # elements from the arena start (multiple of 4)
    numel: int       # padded element count (multiple of 4)
    params: list


def _round_up(n, a):
    return (n + a - 1) // a * a


class ParamArena:
    def __init__(self, params, device, peer=None, bucket_cap_mb=25.0, first_bucket_mb=1.0,
                 local_peers=None, rank=0, last_bucket_mb=4.0):
        """``peer``: PeerComm (multi-process) or None; ``local_peers``: list of PeerComm for the
        single-process multi-GPU engine together with ``rank`` selecting this replica."""
        self.device = torch.device(device)
        self.params = [p for p in params if p.requires_grad]
        self.peer = peer
        order = list(reversed(self.params))
        self.offsets = {}
        self.buckets = []
        cap = int(first_bucket_mb * (1 << 20)) // 4
        off, cur, cur_start = 0, [], 0
        for p in order:
            self.offsets[id(p)] = off
            cur.append(p)
            off += _round_up(p.numel(), ALIGN)
            if off - cur_start >= cap:
                self.buckets.append(Bucket(len(self.buckets), cur_start, off - cur_start, cur))
                cur, cur_start = [], off
                cap = int(bucket_cap_mb * (1 << 20)) // 4
        if cur:
            self.buckets.append(Bucket(len(self.buckets), cur_start, off - cur_start, cur))
        # The last bucket becomes ready only when backward ends, so its collective (and the update and
        # re-pack behind it) has nothing to hide under: keep it small.  Its parameters are the first
        # layers of the model, which are the ones that finish last.
        tail_cap = int(last_bucket_mb * (1 << 20)) // 4
        if self.buckets and last_bucket_mb > 0 and self.buckets[-1].numel > tail_cap and len(self.buckets[-1].params) > 1:
            last = self.buckets.pop()
            tail, tail_n = [], 0
            for p in reversed(last.params):
                n = _round_up(p.numel(), ALIGN)
                if tail and tail_n + n > tail_cap:
                    break
                tail.insert(0, p)
                tail_n += n
            head = last.params[: len(last.params) - len(tail)]
            if head:
                self.buckets.append(Bucket(len(self.buckets), last.offset, last.numel - tail_n, head))
            self.buckets.append(Bucket(len(self.buckets), last.offset + last.numel - tail_n, tail_n, tail))
        self.numel = max(off, ALIGN)
        self.bucket_of = {id(p): b.index for b in self.buckets for p in b.params}
        self.param_buf = self.grad_buf = None
        self._alloc()
        self._adopt()

    # ------------------------------------------------------------------
    def _alloc(self):
        nbytes = self.numel * 4
        if self.peer is not None:
            self.param_buf = self.peer.alloc(nbytes, "params")
            self.grad_buf = self.peer.alloc(nbytes, "grads")
            self.param_flat = self.param_buf.view(torch.float32, 0, self.numel)
            self.grad_flat = self.grad_buf.view(torch.float32, 0, self.numel)
        else:
            self.param_flat = torch.zeros(self.numel, dtype=torch.float32, device=self.device)
            self.grad_flat = torch.zeros(self.numel, dtype=torch.float32, device=self.device)

    def adopt_symmetric(self, param_buf, grad_buf):
        """Single-process engine: arenas were allocated for all replicas at once."""
        self.param_buf, self.grad_buf = param_buf, grad_buf
        old = self.param_flat
        self.param_flat = param_buf.view(torch.float32, 0, self.numel)
        self.param_flat.copy_(old)
        self.grad_flat = grad_buf.view(torch.float32, 0, self.numel)
        self._adopt(copy=False)

    def _adopt(self, copy=True):
        """Re-points every parameter (and its gradient target) into the arenas."""
        with torch.no_grad():
            for p in self.params:
                off = self.offsets[id(p)]
                view = self.param_flat[off: off + p.numel()].view(p.shape)
                if copy:
                    view.copy_(p.data)
                p.data = view
                gview = self.grad_flat[off: off + p.numel()].view(p.shape)
                p._dtdl_grad = gview
                if p.grad is not None:  # adopted mid-iteration: keep what backward produced
                    gview.copy_(p.grad)
                    p.grad = gview
                    p._dtdl_fresh = False
                else:
                    p._dtdl_fresh = True
                p._dtdl_arena = self

    def enable_bf16_grads(self):
        """Second, bf16 gradient arena for the reducer's bf16 transport option (symmetric like the
        fp32 one; the fp32 gradients stay where the kernels write them and are cast per bucket)."""
        if getattr(self, "grad16_buf", None) is None:
            if self.peer is None:
                raise RuntimeError("bf16 gradient transport needs the peer-memory back end")
            self.grad16_buf = self.peer.alloc(self.numel * 2, "grads16")
            self.grad16_flat = self.grad16_buf.view(torch.bfloat16, 0, self.numel)
        return self.grad16_buf

    def flat_view_like(self, flat, p):
        off = self.offsets[id(p)]
        return flat[off: off + p.numel()].view(p.shape)

    def new_state(self, symmetric=False):
        """Zero-initialised optimizer-state arena (momentum / Adam moments)."""
        if symmetric and self.peer is not None:
            buf = self.peer.alloc(self.numel * 4, "opt_state")
            t = buf.view(torch.float32, 0, self.numel)
            t._dtdl_symm = buf
            return t
        return torch.zeros(self.numel, dtype=torch.float32, device=self.device)


class BufferArena:
    """All module buffers (BN running stats, counters) in one byte arena for the broadcast."""

    def __init__(self, buffers, device, peer=None):
        self.device = torch.device(device)
        self.buffers = [b for b in buffers if b is not None]
        self.peer = peer
        off, self.offsets = 0, []
        for b in self.buffers:
            self.offsets.append(off)
            off += _round_up(b.numel() * b.element_size(), 16)
        self.nbytes = max(off, 16)
        if peer is not None:
            self.buf = peer.alloc(self.nbytes, "buffers")
            self.bytes = self.buf.view(torch.uint8, 0, self.nbytes)
        else:
            self.buf = None
            self.bytes = torch.zeros(self.nbytes, dtype=torch.uint8, device=self.device)
        self._adopt()

    def adopt_symmetric(self, buf):
        old = self.bytes
        self.buf = buf
        self.bytes = buf.view(torch.uint8, 0, self.nbytes)
        self.bytes.copy_(old)
        self._adopt(copy=False)

    def _adopt(self, copy=True):
        with torch.no_grad():
            for b, off in zip(self.buffers, self.offsets):
                n = b.numel() * b.element_size()
                view = self.bytes[off: off + n].view(b.dtype).view(b.shape)
                if copy:
                    view.copy_(b.data)
                b.data = view
