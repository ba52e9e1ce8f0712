"""Data-parallel wrapper over the flat gradient arena.

Replaces ``DistributedDataParallel(model, device_ids=[local_rank], output_device=local_rank)``
(imagenet.py:316) and the C++ Reducer behind it (reducer.hpp:347-524):

* construction: parameters and BatchNorm buffers are broadcast from rank 0 (``_sync_module_states``,
  torch/nn/parallel/distributed.py:883-890);
* every training iteration the BatchNorm buffers of rank 0 overwrite everyone's before the next
  forward (``_sync_buffers``, distributed.py:1590-1591,2214-2257).  Here that is ONE broadcast of the
  contiguous buffer arena, issued right after the forward that updated the statistics so it overlaps
  the backward, and joined at the end of that backward (or at the next forward if there is none);
* backward: gradients are written by the model straight into the flat arena (no bucket copy-in /
  copy-out); as the backward passes each bucket boundary the slice is all-reduced (SUM) with NCCL on
  its own stream while the rest of the backward keeps the GPU busy.  The 1/world_size pre-division
  (reducer.hpp:497-500) is folded into the first backward op, so after ``loss.backward()`` ``p.grad``
  holds the same averaged gradient the reference sees;
* ``state_dict()`` keys carry the ``module.`` prefix exactly like the reference's checkpoint
  (imagenet.py:392) and are detached clones in standard (contiguous) layout.
"""
import torch
import torch.distributed as dist
import torch.nn as nn

from . import _nvtx


class FlatDDP(nn.Module):
    def __init__(self, module, device_ids=None, output_device=None, process_group=None, bucket_cap_mb=25,
                 first_bucket_mb=1, broadcast_buffers=True, exchange="nccl", optimizer=None, exchange_ctas=32):
        super().__init__()
        if not hasattr(module, "arena"):
            raise TypeError("FlatDDP wraps a B200ResNet (flat parameter arena); got " + type(module).__name__)
        self.module = module
        self.process_group = process_group
        self.world_size = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.broadcast_buffers = broadcast_buffers
        self._works = []
        self._buffer_work = None
        module.attach_sink(self)
        module.grad_predivide = 1.0 / self.world_size
        a = module.arena
        # buckets: walk the backward's progress marks from the end of the arena, cut when a bucket
        # reaches the cap (first = last layers' bucket is small so communication starts early)
        marks = module.mark_offsets()  # descending, ends with 0
        self.buckets = []
        hi, cap = a.numel, first_bucket_mb * (1 << 20)
        for m in marks:
            if (hi - m) * 4 >= cap or m == 0:
                if hi > m:
                    self.buckets.append((m, hi))
                    hi, cap = m, bucket_cap_mb * (1 << 20)
        self._next = 0
        if self.world_size > 1:
            dist.broadcast(a.params, 0, group=process_group)
            dist.broadcast(module.buffer_arena, 0, group=process_group)
            a.refresh_shadow()
        if exchange not in ("nccl", "peer"):
            raise ValueError("exchange must be 'nccl' or 'peer'")
        self.exchange = exchange if self.world_size > 1 else "nccl"
        self.optimizer = optimizer
        if self.exchange == "peer":
            self._setup_peer(exchange_ctas)

    # ---- opt-in: bucket exchange + SGD as one kernel over NVLink peer memory (K4c, csrc/peer_exchange.cu) ----
    def _setup_peer(self, ctas):
        import torch.distributed._symmetric_memory as symm_mem
        from . import _cabi

        a = self.module.arena
        if not getattr(a, "symmetric", False):
            raise TypeError("exchange='peer' needs build_model(..., symmetric=True): the arenas must be peer-mappable")
        if self.world_size > 8 or len(self.buckets) > 23:
            raise ValueError("exchange='peer' supports up to 8 ranks of one box and 23 buckets")
        group = self.process_group if self.process_group is not None else dist.group.WORLD
        self._peer_rank = dist.get_rank(group)
        self._pad = symm_mem.empty(_cabi.peer_pad_words(), dtype=torch.int32, device=a.params.device).zero_()
        handles = [symm_mem.rendezvous(t, group) for t in (a.grads, a.params, self._pad)]
        hs = symm_mem.rendezvous(a.shadow, group) if a.shadow is not None else None
        self._peer_handles = handles + [hs]   # keep the mappings alive
        self._tab_grad = _cabi.peer_table(handles[0].buffer_ptrs)
        self._tab_param = _cabi.peer_table(handles[1].buffer_ptrs)
        self._tab_pad = _cabi.peer_table(handles[2].buffer_ptrs)
        self._tab_shadow = _cabi.peer_table(hs.buffer_ptrs) if hs is not None else None
        hb = symm_mem.rendezvous(self.module.buffer_arena, group)
        self._peer_handles.append(hb)
        self._tab_buf = _cabi.peer_table(hb.buffer_ptrs)
        self._bcast_issued = False
        self._xstream = torch.cuda.Stream(device=a.params.device)
        self._xctas = int(ctas)
        torch.cuda.synchronize()
        dist.barrier(group=group)          # every pad is zero before anyone raises a flag

    def attach_optimizer(self, optimizer):
        """exchange='peer' applies the optimizer's hyper-parameters inside the backward; FlatSGD.step() then only
        advances the step count."""
        self.optimizer = optimizer

    def _peer_bucket(self, index, lo, hi):
        from . import _cabi

        if self.optimizer is None:
            raise RuntimeError("exchange='peer' needs the optimizer: FlatDDP(..., optimizer=opt) or attach_optimizer(opt)")
        g = self.optimizer.param_groups[0]
        a = self.module.arena
        main = torch.cuda.current_stream()
        self._xstream.wait_stream(main)    # the gradients of this bucket are final in main-stream order
        with torch.cuda.stream(self._xstream):
            _cabi.peer_exchange_sgd(self._tab_grad, self._tab_param, self._tab_shadow, self._tab_pad, a.momentum,
                                    self._peer_rank, self.world_size, index, lo, hi, float(g["lr"]),
                                    float(g["weight_decay"]), float(g["momentum"]), 1.0, a.steps == 0, self._xctas)

    def peer_error(self):
        """True if a kernel of the peer exchange gave up waiting for another rank (call after a synchronise)."""
        return self.exchange == "peer" and bool(self._pad[0].item() != 0)

    # ---- called by the model's backward ---------------------------------------------------------
    def wants(self, offset):
        """True when gradients at arena offsets >= ``offset`` being final completes a bucket."""
        return (self.world_size > 1 and self._next < len(self.buckets) and self.buckets[self._next][0] >= offset)

    def grad_ready(self, offset):
        if self.world_size == 1:
            return
        g = self.module.arena.grads
        while self._next < len(self.buckets) and self.buckets[self._next][0] >= offset:
            lo, hi = self.buckets[self._next]
            _nvtx.push("b200.exchange")
            if self.exchange == "peer":
                self._peer_bucket(self._next, lo, hi)
            else:
                self._works.append(dist.all_reduce(g[lo:hi], op=dist.ReduceOp.SUM, group=self.process_group,
                                                   async_op=True))
            _nvtx.pop()
            self._next += 1

    def backward_done(self):
        for w in self._works:
            w.wait()  # stream-level wait: the optimizer's kernels queue behind the all-reduces
        self._works = []
        if self.exchange == "peer" and self._next > 0:
            from . import _cabi

            a = self.module.arena
            torch.cuda.current_stream().wait_stream(self._xstream)
            # every shard of every rank has landed; rank 0 also waits until its buffers have been read
            _cabi.peer_wait(self._pad, self.world_size, len(self.buckets), self._peer_rank,
                            self._bcast_issued and self._peer_rank == 0)
            self._bcast_issued = False
            a.applied_step = a.steps                                         # optimizer.step() only counts
        self._next = 0
        if self._buffer_work is not None:
            # the buffer broadcast had the whole backward to hide behind; joining it here (rather than at the
            # next forward) keeps every collective of an iteration inside that iteration - which is also what
            # lets the iteration be captured as one CUDA graph
            self._buffer_work.wait()
            self._buffer_work = None

    def drain(self):
        """Wait (on the stream) for every collective this wrapper has in flight."""
        if self._buffer_work is not None:
            self._buffer_work.wait()
            self._buffer_work = None
        for w in self._works:
            w.wait()
        self._works = []

    # ---- nn.Module surface ----------------------------------------------------------------------
    def forward(self, *args, **kwargs):
        if self._buffer_work is not None:
            self._buffer_work.wait()
            self._buffer_work = None
        out = self.module(*args, **kwargs)
        if (self.world_size > 1 and self.broadcast_buffers and self.module.training and torch.is_grad_enabled()):
            if self.exchange == "peer":
                # the same broadcast as a peer copy (a kernel, so the iteration stays capturable); joined, like the
                # bucket kernels, by backward_done
                from . import _cabi

                self._xstream.wait_stream(torch.cuda.current_stream())
                with torch.cuda.stream(self._xstream):
                    _cabi.peer_bcast(self._tab_buf, self._tab_pad, self.module.buffer_arena.numel(), self._peer_rank,
                                     self.world_size, 0)
                self._bcast_issued = True
            else:
                self._buffer_work = dist.broadcast(self.module.buffer_arena, 0, group=self.process_group,
                                                   async_op=True)
        return out

    def state_dict(self, *args, **kwargs):
        if self._buffer_work is not None:
            self._buffer_work.wait()
            self._buffer_work = None
        sd = super().state_dict(*args, **kwargs)
        return type(sd)((k, v.detach().clone(memory_format=torch.contiguous_format)) for k, v in sd.items())


# the name the reference imports (imagenet.py:28)
DistributedDataParallel = FlatDDP
