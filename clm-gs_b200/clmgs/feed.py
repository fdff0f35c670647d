"""Camera / ground-truth feed: the step immediately before the hot path.

The reference uploads each batch synchronously right before the iteration (train.py:273-312:
``world_view_transform.cuda()``, ``create_k_on_gpu()``, a batched ``torch.inverse`` for
``camtoworlds``, ``original_image_backup.cuda()`` -- 24.9 MB per camera at 4K).  Here the next
batch is staged while the current one trains:

* one pinned staging block per batch slot holds every camera's w2c (row-major), its transpose
  (the reference's ``world_view_transform``), c2w, K and camera position = 64 floats per camera, so
  the matrices of a batch travel in ONE copy; c2w is computed on the host in float64 (4x4);
* the uint8 images are copied from pinned host memory image by image on the feed stream;
* two slots (double buffering): ``get(b)`` makes the consumer's stream wait on slot b's upload
  event, then immediately issues the upload of batch b+1 into the other slot.  A slot is only
  overwritten after the consumer recorded that it finished with it (``release`` event).
"""
import numpy as np
import torch

from .model import Camera

_FLOATS_PER_CAMERA = 16 + 16 + 16 + 9 + 3  # w2c, w2c^T, c2w, K, campos  (= 60, padded to 64; matrices 64-B aligned)
_STRIDE = 64


class _DeviceCamera(Camera):
    """Camera whose tensors are views into a feed slot (no per-field host->device copies)."""

    def __init__(self, block, image, uid):  # noqa: super().__init__ deliberately not called
        self.viewmat = block[0:16].view(4, 4)
        self.world_view_transform = block[16:32].view(4, 4)
        self.camtoworlds = block[32:48].view(1, 4, 4)
        self.K = block[48:57].view(3, 3)
        self.campos = block[57:60]
        self.original_image = image
        self.image_height, self.image_width = int(image.shape[1]), int(image.shape[2])
        self.uid = uid


class CameraFeed:
    def __init__(self, viewmats, Ks, images_u8, bsz, device="cuda:0", image_index=None, stream=None):
        """viewmats [M,4,4], Ks [M,3,3] (numpy or CPU tensors); images_u8 [n_img,3,H,W] uint8 CPU tensor (pinned
        here if it is not); image_index[j] = which image camera j uses (default j % n_img)."""
        self.device = torch.device(device)
        self.bsz = bsz
        V = np.asarray(viewmats, np.float64).reshape(-1, 4, 4)
        K = np.asarray(Ks, np.float64).reshape(-1, 3, 3)
        M = V.shape[0]
        self.n_batches = M // bsz
        c2w = np.linalg.inv(V)
        host = np.zeros((M, _STRIDE), np.float32)
        host[:, 0:16] = V.reshape(M, 16)
        host[:, 16:32] = V.transpose(0, 2, 1).reshape(M, 16)
        host[:, 32:48] = c2w.reshape(M, 16)
        host[:, 48:57] = K.reshape(M, 9)
        host[:, 57:60] = c2w[:, :3, 3]
        self._host = torch.from_numpy(host).pin_memory()
        img = torch.as_tensor(images_u8)
        assert img.dtype == torch.uint8 and img.dim() == 4 and img.shape[1] == 3
        self._images = img if img.is_pinned() else img.pin_memory()
        n_img = img.shape[0]
        self._image_index = (np.arange(M) % n_img) if image_index is None else np.asarray(image_index)
        self.stream = stream or torch.cuda.Stream(self.device)
        self._slots = []
        for _ in range(2):
            self._slots.append(dict(
                block=torch.empty((bsz, _STRIDE), dtype=torch.float32, device=self.device),
                images=torch.empty((bsz,) + tuple(img.shape[1:]), dtype=torch.uint8, device=self.device),
                ready=torch.cuda.Event(), released=None, batch=-1))
        self.h2d_bytes_per_batch = bsz * (_STRIDE * 4 + int(np.prod(img.shape[1:])))

    def prefetch(self, b):
        """Issue the upload of batch b (no-op if it is already staged)."""
        slot = self._slots[b % 2]
        if slot["batch"] == b:
            return
        with torch.cuda.stream(self.stream):
            if slot["released"] is not None:
                self.stream.wait_event(slot["released"])
            lo = (b % self.n_batches) * self.bsz  # batch indices wrap around the camera list (epochs)
            slot["block"].copy_(self._host[lo:lo + self.bsz], non_blocking=True)
            for i in range(self.bsz):
                slot["images"][i].copy_(self._images[int(self._image_index[lo + i])], non_blocking=True)
            slot["ready"].record(self.stream)
        slot["batch"] = b

    def get(self, b):
        """Cameras of batch b, valid on the current stream; starts uploading batch b+1."""
        self.prefetch(b)
        slot = self._slots[b % 2]
        torch.cuda.current_stream(self.device).wait_event(slot["ready"])
        cams = [_DeviceCamera(slot["block"][i], slot["images"][i], (b % self.n_batches) * self.bsz + i)
                for i in range(self.bsz)]
        other = self._slots[(b + 1) % 2]
        if other["batch"] != b + 1:
            # the other slot was last used by batch b-1, whose iteration has been issued completely
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            other["released"] = ev
            self.prefetch(b + 1)
        return cams
