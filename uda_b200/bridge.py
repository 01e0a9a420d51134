"""Python face of the host runtime (``libuda.so``): plays the role of the Java plug-ins.

The reference is driven by ``UdaBridge.java`` through JNI (plugins/shared/com/mellanox/hadoop/mapred/
UdaBridge.java:38-145). There is no JVM on this image, so the same native roles are reachable through a
C API (include/uda_rt_api.h) and this module is the "fake Hadoop": it builds the argv
(``UdaPluginRT.buildCmdParams``, UdaPlugin.java:181-201), forms commands (``UdaCmd.formCmd``,
:577-586), answers the ``getPathUda`` / ``getConfData`` up-calls and collects ``dataFromUda`` blocks.
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Callable, Dict, List, Optional, Sequence, Tuple

from ._lib import runtime

EXIT, NEW_MAP, FINAL_MERGE, RESULT, FETCH, FETCH_OVER, JOB_OVER, INIT, MORE, REDUCE_LAUNCHED = range(10)

_FETCH_OVER = C.CFUNCTYPE(None, C.c_void_p)
_DATA = C.CFUNCTYPE(None, C.c_void_p, C.c_void_p, C.c_int)
_PATH = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_char_p, C.c_char_p, C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong),
                    C.POINTER(C.c_longlong), C.c_void_p, C.c_int)
_CONF = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_char_p, C.c_char_p, C.c_void_p, C.c_int)
_LOG = C.CFUNCTYPE(None, C.c_void_p, C.c_char_p, C.c_int)
_FAIL = C.CFUNCTYPE(None, C.c_void_p)


class _Callbacks(C.Structure):
    _fields_ = [("user", C.c_void_p), ("fetch_over_message", _FETCH_OVER), ("data_from_uda", _DATA), ("get_path_uda", _PATH),
                ("get_conf_data", _CONF), ("log_to_java", _LOG), ("failure_in_uda", _FAIL)]


_bound = False


def lib() -> C.CDLL:
    global _bound
    L = runtime()
    if not _bound:
        L.uda_rt_start.restype = C.c_void_p
        L.uda_rt_start.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_char_p), C.c_int, C.c_int, C.POINTER(_Callbacks)]
        L.uda_rt_do_command.argtypes = [C.c_void_p, C.c_char_p]
        L.uda_rt_reduce_exit.argtypes = [C.c_void_p]
        L.uda_rt_failed.argtypes = [C.c_void_p]
        L.uda_rt_last_error.argtypes = [C.c_void_p]
        L.uda_rt_last_error.restype = C.c_char_p
        L.uda_rt_port.argtypes = [C.c_void_p]
        L.uda_rt_stat.argtypes = [C.c_void_p, C.c_char_p]
        L.uda_rt_stat.restype = C.c_longlong
        L.uda_rt_destroy.argtypes = [C.c_void_p]
        L.uda_rt_version.restype = C.c_char_p
        L.uda_rt_vint_encode.argtypes = [C.c_longlong, C.c_char_p]
        L.uda_rt_vint_decode.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_longlong)]
        L.uda_rt_vint_size.argtypes = [C.c_longlong]
        L.uda_rt_parse_cmd.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.uda_rt_form_cmd.argtypes = [C.c_int, C.c_int, C.POINTER(C.c_char_p), C.c_char_p, C.c_int]
        L.uda_rt_parse_options.argtypes = [C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_longlong), C.c_char_p, C.c_int]
        L.uda_rt_fetch_request_roundtrip.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.uda_rt_fetch_ack_roundtrip.argtypes = [C.c_char_p, C.c_char_p, C.c_int]
        L.uda_rt_compare_keys.argtypes = [C.c_int, C.c_char_p, C.c_int, C.c_char_p, C.c_int]
        L.uda_rt_key_kind.argtypes = [C.c_char_p]
        L.uda_rt_snappy_uncompress.argtypes = [C.c_char_p, C.c_longlong, C.c_char_p, C.c_longlong]
        L.uda_rt_snappy_store.argtypes = [C.c_char_p, C.c_longlong, C.c_char_p, C.c_longlong]
        L.uda_rt_snappy_store.restype = C.c_longlong
        L.uda_rt_plan_memory.argtypes = [C.c_int, C.c_longlong, C.c_longlong, C.c_longlong, C.c_int, C.c_double, C.POINTER(C.c_longlong)]
        _bound = True
    return L


def form_cmd(cmd: int, params: Sequence[str]) -> str:
    """``UdaCmd.formCmd``: ``"<n+1>:<cmd>:<p1>:...:<pn>"``."""
    return ":".join([str(len(params) + 1), str(cmd), *params])


def build_argv(wqes: int = 256, port: int = 9011, approach: int = 1, log_dir: str = "/tmp/", buf_kb: int = 1024,
               provider: bool = False) -> List[str]:
    """The argv Java hands to ``startNative`` (no program name; the provider omits ``-a``)."""
    a = ["-w", str(wqes), "-r", str(port)]
    if not provider:
        a += ["-a", str(approach)]
    a += ["-m", "1", "-g", log_dir, "-s", str(buf_kb)]
    return a


class NativeRole:
    """One native role (provider or consumer) with Python callables as the Java up-call targets."""

    def __init__(self, is_net_merger: bool, argv: Sequence[str], log_level: int = 4, log_to_unique_file: bool = False,
                 get_path: Optional[Callable[[str, str, int], Optional[Tuple[int, int, int, str]]]] = None,
                 conf: Optional[Dict[str, str]] = None):
        self.blocks: List[bytes] = []
        self.fetch_over = 0
        self.failures = 0
        self.logs: List[Tuple[int, str]] = []
        self.conf = dict(conf or {})
        self._get_path = get_path
        self._lock = threading.Lock()
        self._cb = _Callbacks(
            None, _FETCH_OVER(self._on_fetch_over), _DATA(self._on_data), _PATH(self._on_path), _CONF(self._on_conf),
            _LOG(self._on_log), _FAIL(self._on_fail))
        arr = (C.c_char_p * len(argv))(*[a.encode() for a in argv])
        self.h = lib().uda_rt_start(int(is_net_merger), len(argv), arr, log_level, int(log_to_unique_file), C.byref(self._cb))

    # ---- up-calls -------------------------------------------------------------------------------
    def _on_fetch_over(self, _u):
        with self._lock:
            self.fetch_over += 1

    def _on_data(self, _u, buf, n):
        with self._lock:
            self.blocks.append(C.string_at(buf, n))

    def _on_path(self, _u, job, mapid, reduce, off, raw, part, path_out, cap):
        if self._get_path is None:
            return 0
        r = self._get_path(job.decode(), mapid.decode(), reduce)
        if r is None:
            return 0
        off[0], raw[0], part[0] = r[0], r[1], r[2]
        p = r[3].encode()[: cap - 1]
        C.memmove(path_out, p + b"\0", len(p) + 1)
        return 1

    def _on_conf(self, _u, name, dflt, out, cap):
        v = self.conf.get(name.decode())
        if v is None:
            return 0
        b = v.encode()[: cap - 1]
        C.memmove(out, b + b"\0", len(b) + 1)
        return 1

    def _on_log(self, _u, msg, sev):
        with self._lock:
            self.logs.append((sev, msg.decode(errors="replace")))

    def _on_fail(self, _u):
        with self._lock:
            self.failures += 1

    # ---- down-calls -----------------------------------------------------------------------------
    def command(self, cmd: str) -> int:
        return lib().uda_rt_do_command(self.h, cmd.encode())

    def reduce_exit(self) -> None:
        lib().uda_rt_reduce_exit(self.h)

    @property
    def failed(self) -> bool:
        return bool(lib().uda_rt_failed(self.h))

    @property
    def last_error(self) -> str:
        return lib().uda_rt_last_error(self.h).decode(errors="replace")

    @property
    def port(self) -> int:
        return lib().uda_rt_port(self.h)

    def stat(self, what: str) -> int:
        return int(lib().uda_rt_stat(self.h, what.encode()))

    def stream(self) -> bytes:
        with self._lock:
            return b"".join(self.blocks)

    def close(self) -> None:
        if self.h:
            lib().uda_rt_destroy(self.h)
            self.h = None
