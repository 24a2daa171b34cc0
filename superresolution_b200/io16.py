"""16-bit edges and preview glue of the reference's callers on libsrb:
``save_as_tiff16`` (scripts/Modello_5.py:29-33), the TIFF -> [0,1] tensor load of ``src/dataset.py:24-41`` and the
[LR | SR | HR] comparison strip (scripts/Modello_5.py:83-85, scripts/Modello_supporto.py:187-190).

The conversions run on the device (2 bytes per pixel cross PCIe instead of 4); the TIFF container itself is written / read
on the host by a ~40-line baseline codec (little-endian, one uncompressed strip, 16-bit grayscale — what PIL's
``Image.fromarray(arr, mode='I;16').save(path)`` produces and reads back)."""
import struct

import numpy as np
import torch

from . import _lib as L


def to_u16(t):
    """u16(clip(t, 0, 1) * 65535) with a truncating cast (Modello_5.py:31-32) — fp32 CUDA tensor -> uint16 CUDA tensor."""
    if t.device.type != 'cuda':
        raise RuntimeError('to_u16 (libsrb) runs on CUDA only; no CPU fallback')
    x = t.float().contiguous()
    out = torch.empty(x.shape, dtype=torch.uint16, device=x.device)
    with torch.cuda.device(x.device):
        L.check(L.lib().srb_to_u16(L.ptr(x), L.ptr(out), x.numel(), L.current_stream()), 'to_u16')
    return out


def write_tiff16(arr_u16, path):
    """[H,W] uint16 numpy array -> baseline TIFF (II, one strip, no compression, BlackIsZero)."""
    a = np.ascontiguousarray(arr_u16, dtype='<u2')
    if a.ndim != 2:
        raise ValueError(f'write_tiff16: [H,W] expected, got {a.shape}')
    h, w = a.shape
    tags = [(256, 4, 1, w), (257, 4, 1, h), (258, 3, 1, 16), (259, 3, 1, 1), (262, 3, 1, 1), (273, 4, 1, 8),
            (277, 3, 1, 1), (278, 4, 1, h), (279, 4, 1, a.nbytes)]
    ifd_off = 8 + a.nbytes + (a.nbytes & 1)
    with open(path, 'wb') as f:
        f.write(struct.pack('<2sHI', b'II', 42, ifd_off))
        f.write(a.tobytes())
        if a.nbytes & 1:
            f.write(b'\0')
        f.write(struct.pack('<H', len(tags)))
        for tag, typ, cnt, val in tags:
            f.write(struct.pack('<HHI', tag, typ, cnt) + (struct.pack('<HH', val, 0) if typ == 3 else struct.pack('<I', val)))
        f.write(struct.pack('<I', 0))


def read_tiff16(path):
    """baseline 16-bit grayscale TIFF (uncompressed, either byte order, any strip layout) -> [H,W] uint16 numpy array."""
    with open(path, 'rb') as f:
        buf = f.read()
    bo = {b'II': '<', b'MM': '>'}.get(buf[:2])
    if bo is None or struct.unpack(bo + 'H', buf[2:4])[0] != 42:
        raise ValueError(f'{path}: not a TIFF file')
    off = struct.unpack(bo + 'I', buf[4:8])[0]
    n = struct.unpack(bo + 'H', buf[off:off + 2])[0]
    size = {1: 1, 3: 2, 4: 4}
    tags = {}
    for i in range(n):
        tag, typ, cnt = struct.unpack(bo + 'HHI', buf[off + 2 + 12 * i:off + 10 + 12 * i])
        raw = buf[off + 10 + 12 * i:off + 14 + 12 * i]
        if typ not in size:
            continue
        if size[typ] * cnt > 4:
            p = struct.unpack(bo + 'I', raw)[0]
            raw = buf[p:p + size[typ] * cnt]
        tags[tag] = list(struct.unpack(bo + {1: 'B', 3: 'H', 4: 'I'}[typ] * cnt, raw[:size[typ] * cnt]))
    w, h = tags[256][0], tags[257][0]
    if tags.get(258, [1])[0] != 16 or tags.get(259, [1])[0] != 1 or tags.get(277, [1])[0] != 1:
        raise ValueError(f'{path}: only uncompressed single-channel 16-bit TIFF is supported')
    data = b''.join(buf[o:o + c] for o, c in zip(tags[273], tags.get(279, [w * h * 2])))
    return np.frombuffer(data, dtype=bo + 'u2', count=w * h).reshape(h, w).astype(np.uint16)


def save_as_tiff16(tensor, path):
    """scripts/Modello_5.py:29-33: squeeze, clip to [0,1], * 65535, truncate to uint16, write a 16-bit TIFF."""
    u16 = to_u16(tensor.squeeze())
    write_tiff16(u16.cpu().numpy(), path)


def load_tiff_as_tensor(path, device='cuda'):
    """src/dataset.py:24-41 without the float round trip on the host: the uint16 counts go to the device as they are
    ([1,H,W] uint16); ``HybridSuperResolutionModel`` takes them directly (the division by 65535 is fused into its first
    kernel), ``.float() / 65535`` gives the reference's tensor."""
    return torch.from_numpy(read_tiff16(path)).unsqueeze(0).to(device)


def preview_strip(lr, sr, hr, quantize=False):
    """cat(F.interpolate(lr, size=sr.shape[-2:], mode='nearest'), sr, hr, dim=3).clamp(0, 1)  ([B,1,H,3W] fp32), or with
    ``quantize`` the uint8 image torchvision's save_image writes (x * 255 + 0.5, clamp, truncate)."""
    if sr.device.type != 'cuda':
        raise RuntimeError('preview_strip (libsrb) runs on CUDA only; no CPU fallback')
    lr, sr, hr = (t.float().contiguous() for t in (lr, sr, hr))
    B, _, H, W = sr.shape
    if hr.shape != sr.shape or lr.shape[0] != B or lr.shape[1] != 1 or sr.shape[1] != 1:
        raise ValueError('preview_strip: lr [B,1,h,w], sr and hr [B,1,H,W] expected')
    out = torch.empty(B, 1, H, 3 * W, dtype=torch.uint8 if quantize else torch.float32, device=sr.device)
    with torch.cuda.device(sr.device):
        L.check(L.lib().srb_preview_strip(L.ptr(lr), L.ptr(sr), L.ptr(hr), L.ptr(out), int(quantize), B, lr.shape[2],
                                          lr.shape[3], H, W, L.current_stream()), 'preview_strip')
    return out
