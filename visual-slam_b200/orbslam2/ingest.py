"""Video ingest on the device: the reference's `cv2.VideoCapture(path).read()` loop
(src/tests/tester_map.py:58-66, run_video.py:138-149) with the decode and the colour conversion on the GPU.

`VideoReader(path)` demuxes the container on the host -- cv2.VideoCapture in raw mode (CAP_PROP_FORMAT = -1) hands
out the encoded packets; container parsing is IO, not compute -- and feeds them to NVDEC through
libvslam_b200.so (csrc/nvdec.cu).  Frames come back as torch uint8 tensors in HBM, ready for
vslam_orb_detect_and_compute / vslam_frontend_run without a host hop:

    mode "gray"  the reference's gray frame bit for bit (swscale yuv420p -> bgr24, then cvtColor(BGR2GRAY))
    mode "bgr"   what cv2.VideoCapture.read() returns
    mode "luma"  the decoder's Y plane (native, not the reference's gray)
    mode "nv12"  the decoder's surface as is

There is no CPU fallback: without libnvcuvid (the NVDEC driver library) or a GPU the reader raises.
"""
import ctypes as C

import numpy as np

from . import _native

_MODES = {"luma": 0, "gray": 1, "bgr": 2, "nv12": 3}
# FFmpeg codec tags as cv2 reports them (CAP_PROP_FOURCC) -> VSLAM_CODEC_*
_CODECS = {"mp4v": 1, "FMP4": 1, "MP4V": 1, "avc1": 2, "h264": 2, "H264": 2, "MJPG": 3, "mjpg": 3, "hev1": 4, "hvc1": 4,
           "HEVC": 4, "VP80": 5, "VP90": 6, "vp09": 6, "av01": 7, "AV01": 7}


def _bind(L):
    vp, i32 = C.c_void_p, C.c_int
    L.vslam_nvdec_create.argtypes = [i32, C.POINTER(vp)]
    L.vslam_nvdec_destroy.argtypes = [vp]
    L.vslam_nvdec_set_colour.argtypes = [vp, i32, i32, i32, i32, i32, i32]
    L.vslam_nvdec_frame_size.argtypes = [vp, vp, vp]
    L.vslam_nvdec_decode.argtypes = [vp, vp, C.c_size_t, i32, i32, vp, i32, C.c_size_t, i32, vp, vp]
    return L


def available():
    """True when the NVDEC driver library is loadable (vslam_nvdec_available)."""
    return bool(_native.lib().vslam_nvdec_available())


class VideoReader:
    """Iterates the frames of a video file as device tensors.  `mode` as in the module docstring."""

    def __init__(self, path, mode="gray", max_frames_per_packet=8):
        import cv2
        import torch
        _native.require_gpu()
        self._L = _bind(_native.lib())
        if mode not in _MODES:
            raise ValueError("mode must be one of %s" % sorted(_MODES))
        self.mode, self._m = mode, _MODES[mode]
        self._cap = cv2.VideoCapture(path, cv2.CAP_FFMPEG, [cv2.CAP_PROP_FORMAT, -1])      # raw mode: encoded packets
        if not self._cap.isOpened():
            raise IOError("cannot open %r" % (path,))
        fourcc = int(self._cap.get(cv2.CAP_PROP_FOURCC)).to_bytes(4, "little").decode("latin1")
        if fourcc not in _CODECS:
            raise NotImplementedError("codec %r is not routed to NVDEC" % fourcc)
        self.codec = fourcc
        self.width = int(self._cap.get(cv2.CAP_PROP_FRAME_WIDTH))
        self.height = int(self._cap.get(cv2.CAP_PROP_FRAME_HEIGHT))
        self._h = C.c_void_p()
        _native.check(self._L.vslam_nvdec_create(_CODECS[fourcc], C.byref(self._h)))
        self._torch = torch
        self._cap_frames = int(max_frames_per_packet)
        self._stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        self._eos = False
        # sequence headers that the container keeps outside the packets (MPEG-4 VOL, H.264 SPS / PPS in avcC form
        # are converted by FFmpeg's bitstream filters in raw mode); fed first when present
        self._extradata = None
        idx = int(self._cap.get(cv2.CAP_PROP_CODEC_EXTRADATA_INDEX))
        if idx >= 0:
            ok, ed = self._cap.retrieve(flag=idx)
            if ok and ed is not None and ed.size:
                self._extradata = np.ascontiguousarray(ed).ravel()

    def set_colour(self, cy, cbu, cgu, cgv, crv, y_offset):
        _native.check(self._L.vslam_nvdec_set_colour(self._h, cy, cbu, cgu, cgv, crv, y_offset))

    def _shape(self):
        if self.mode == "bgr":
            return (self._cap_frames, self.height, self.width, 3)
        if self.mode == "nv12":
            return (self._cap_frames, self.height + self.height // 2, self.width)
        return (self._cap_frames, self.height, self.width)

    def _decode(self, payload, eos):
        torch = self._torch
        out = torch.empty(self._shape(), dtype=torch.uint8, device="cuda")
        n = C.c_int()
        pitch = out.stride(1)
        p = None if payload is None else payload.ctypes.data_as(C.c_void_p)
        _native.check(self._L.vslam_nvdec_decode(self._h, p, 0 if payload is None else payload.size, int(eos), self._m,
                                                 C.c_void_p(out.data_ptr()), pitch, out.stride(0), self._cap_frames,
                                                 C.byref(n), self._stream))
        return [out[i] for i in range(n.value)]

    def __iter__(self):
        if self._extradata is not None:
            for f in self._decode(self._extradata, False):
                yield f
        while True:
            if not self._cap.grab():
                break
            ok, pkt = self._cap.retrieve()
            if not ok or pkt is None:
                break
            for f in self._decode(np.ascontiguousarray(pkt).ravel(), False):
                yield f
        for f in self._decode(None, True):          # flush the frames the decoder still holds
            yield f

    def close(self):
        if self._h:
            self._L.vslam_nvdec_destroy(self._h)
            self._h = C.c_void_p()
        if self._cap is not None:
            self._cap.release()
            self._cap = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
