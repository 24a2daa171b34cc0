"""Host-to-host inference over a stream of batches — the loop of the reference's inference script
(scripts/Modello_5.py:35-91 moves every patch batch to the GPU, runs the model, brings the result back), with the
copies taken off the critical path: the H2D copy of batch i+1 and the D2H copy of batch i-1 run on their own streams
under the forward of batch i (two device input buffers; the output of a forward is handed to the copy-out stream).
Inputs and outputs should be pinned (``tensor.pin_memory()``) for the copies to be asynchronous."""
import torch


def infer_host_batches(model, batches, out=None):
    """batches: iterable of host tensors [B,1,H,W] (float32 in [0,1] or uint16), all of one shape and dtype.
    out: optional list of host tensors [B,1,4H,4W] to receive the results (may repeat one buffer; allocated pinned
    when omitted).  Returns the list of output host tensors; they are complete when the call returns."""
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
        x_dev = [torch.empty(batches[0].shape, dtype=batches[0].dtype, device=dev) for _ in range(min(2, n))]
        ready = [torch.cuda.Event() for _ in range(n)]
        done = [torch.cuda.Event() for _ in range(n)]
        copied = [torch.cuda.Event() for _ in range(n)]
        H, W = batches[0].shape[-2:]
        direct = hasattr(model, 'run') and 4 * W == getattr(model, 'output_size', -1)  # no resize branch: write in place
        y_dev = [None, None]  # two device result buffers in rotation (a fresh tensor per forward would have to outlive
        results = []          # its copy on another stream: the allocator could not recycle it in time)

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
            if direct:
                if y_dev[i % 2] is None:
                    y_dev[i % 2] = torch.empty(batches[0].shape[0], 1, 4 * H, 4 * W, dtype=torch.float32, device=dev)
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
