"""Host-to-host inference over a stream of batches — the loop of the reference's inference script
(scripts/Modello_5.py:35-91 moves every patch batch to the GPU, runs the model, brings the result back), with the
copies taken off the critical path: the H2D copy of batch i+1 and the D2H copy of batch i-1 run on their own streams
under the forward of batch i (two device input buffers; the output of a forward is handed to the copy-out stream).
Inputs and outputs should be pinned (``tensor.pin_memory()``) for the copies to be asynchronous."""
import torch


def _slots(model, batches, dev, use_graphs):
    """Device input / result buffers of the two pipeline slots (kept on the model between calls) and, when the forward
    can be replayed, one CUDA graph per slot.  A graph is tied to its buffers, to the packed weights and to the
    workspace: any of them moving invalidates the cached entry."""
    B, _, H, W = batches[0].shape
    direct = hasattr(model, 'run') and 4 * W == getattr(model, 'output_size', -1)  # no resize branch: result in place
    packed = model.pack_weights() if direct else None
    ws = model._workspace(B, H, W, dev) if direct else None
    key = (tuple(batches[0].shape), batches[0].dtype, dev, None if packed is None else packed.data_ptr(),
           None if ws is None else ws.data_ptr(), bool(use_graphs and direct))
    cache = model.__dict__.setdefault('_pipe_slots', {})
    ent = cache.get('slots')
    if ent is not None and ent['key'] == key:
        return ent
    x_dev = [torch.empty(batches[0].shape, dtype=batches[0].dtype, device=dev) for _ in range(2)]
    y_dev = [torch.empty(B, 1, 4 * H, 4 * W, dtype=torch.float32, device=dev) if direct else None for _ in range(2)]
    graphs = [None, None]
    if use_graphs and direct:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for k in range(2):
                x_dev[k].zero_()
                model.run(x_dev[k], 3, out=y_dev[k])  # warm-up outside the capture
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize(dev)
        for k in range(2):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                model.run(x_dev[k], 3, out=y_dev[k])
            graphs[k] = g
    ent = {'key': key, 'x': x_dev, 'y': y_dev, 'graphs': graphs, 'direct': direct}
    cache['slots'] = ent
    return ent


def infer_host_batches(model, batches, out=None, use_graphs=True):
    """batches: iterable of host tensors [B,1,H,W] (float32 in [0,1] or uint16), all of one shape and dtype.
    out: optional list of host tensors [B,1,4H,4W] to receive the results (may repeat one buffer; allocated pinned
    when omitted).  use_graphs: replay the forward of each slot as a CUDA graph (captured once per model, shape and
    weight buffer; eval-mode full-model forwards only).  Returns the list of output host tensors; they are complete
    when the call returns."""
    batches = list(batches)
    if not batches:
        return []
    dev = next(model.parameters()).device
    if dev.type != 'cuda':
        raise RuntimeError('infer_host_batches: the model must live on a CUDA device (libsrb has no CPU path)')
    n = len(batches)
    if out is not None and len(out) != n:
        raise ValueError('infer_host_batches: len(out) must equal the number of batches')
    with torch.cuda.device(dev), torch.no_grad():
        cur = torch.cuda.current_stream()
        s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
        s_in.wait_stream(cur)
        s_out.wait_stream(cur)
        slots = _slots(model, batches, dev, use_graphs)  # (captures the two forward graphs on first use)
        x_dev, y_dev, graphs, direct = slots['x'], slots['y'], slots['graphs'], slots['direct']
        ready = [torch.cuda.Event() for _ in range(n)]
        done = [torch.cuda.Event() for _ in range(n)]
        copied = [torch.cuda.Event() for _ in range(n)]
        results = []  # (two device result buffers in rotation: a fresh tensor per forward would have to outlive its copy
                      # on another stream, and the allocator could not recycle it in time)

        def load(i):
            with torch.cuda.stream(s_in):
                if i >= 2:
                    s_in.wait_event(done[i - 2])  # the forward that read this buffer has finished
                x_dev[i % 2].copy_(batches[i], non_blocking=True)
                ready[i].record(s_in)

        load(0)
        for i in range(n):
            if i + 1 < n:
                load(i + 1)
            cur.wait_event(ready[i])
            if i >= 2:
                cur.wait_event(copied[i - 2])  # the result buffer of this slot has left the device
            if graphs[i % 2] is not None:
                graphs[i % 2].replay()
                y = y_dev[i % 2]
            elif direct:
                y = model.run(x_dev[i % 2], 3, out=y_dev[i % 2])
            else:
                y = model(x_dev[i % 2])
                if y_dev[i % 2] is None:
                    y_dev[i % 2] = torch.empty_like(y)
                y_dev[i % 2].copy_(y)
                y = y_dev[i % 2]
            done[i].record(cur)
            with torch.cuda.stream(s_out):
                s_out.wait_event(done[i])
                h = torch.empty(y.shape, dtype=y.dtype).pin_memory() if out is None else out[i]
                h.copy_(y, non_blocking=True)
                copied[i].record(s_out)
                results.append(h)
        cur.wait_stream(s_out)
        cur.wait_stream(s_in)
    torch.cuda.current_stream(dev).synchronize()
    return results
