"""One typed configuration object that accepts the reference's key names.

The reference reads its tunables in three tiers: Hadoop JobConf keys in Java (plugins/shared/.../UdaPlugin.java:
185-293), a getopt argv handed to ``startNative`` (src/CommUtils/C2JNexus.cc:43-137) and late ``getConfData``
look-ups from C++ (src/UdaBridge.cc:419-438). :class:`UdaConfig` holds all of them under their Hadoop names, builds
the same argv / INIT parameter lists the Java side would build, answers ``getConfData`` for the fake-Hadoop driver,
and carries the B200-only knobs (HBM budget, sub-partitions, push CTAs) under a ``uda.b200.`` prefix.
"""
from __future__ import annotations

import dataclasses
import math
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import Dict, List, Mapping, Optional, Sequence

#: field name -> (Hadoop key, type)
_KEYS = {
    "wqe_per_conn": ("mapred.rdma.wqe.per.conn", int),
    "cma_port": ("mapred.rdma.cma.port", int),
    "merge_approach": ("mapred.netmerger.merge.approach", int),
    "buf_size_kb": ("mapred.rdma.buf.size", int),
    "buf_size_min_kb": ("mapred.rdma.buf.size.min", int),
    "shuffle_total_size": ("mapred.rdma.shuffle.total.size", int),
    "shuffle_input_buffer_percent": ("mapred.job.shuffle.input.buffer.percent", float),
    "hybrid_lpq_size": ("mapred.netmerger.hybrid.lpq.size", int),
    "num_parallel_lpqs": ("mapred.rdma.num.parallel.lpqs", int),
    "compression_buffer_ratio": ("mapred.rdma.compression.buffer.ratio", float),
    "mem_use_contig_pages": ("mapred.rdma.mem.use.contig.pages", int),
    "map_output_compression_codec": ("mapred.map.output.compression.codec", str),
    "compress_map_output": ("mapred.compress.map.output", bool),
    "lzo_buffersize": ("io.compression.codec.lzo.buffersize", int),
    "snappy_buffersize": ("io.compression.codec.snappy.buffersize", int),
    "lzo_decompressor": ("io.compression.codec.lzo.decompressor", str),
    "log_to_unique_file": ("mapred.uda.log.to.unique.file", bool),
    "developer_mode": ("mapred.rdma.developer.mode", bool),
    "blocked_threads_per_disk": ("mapred.uda.provider.blocked.threads.per.disk", int),
    "provider_reader": ("mapred.uda.provider.reader", str),
    "decompress_ahead": ("mapred.uda.decompress.ahead", bool),
    "local_dirs": ("mapred.local.dir", list),
    "map_output_key_class": ("mapred.mapoutput.key.class", str),
    "child_java_opts": ("mapred.child.java.opts", str),
    # B200-only
    "hbm_budget_bytes": ("uda.b200.hbm.budget.bytes", int),
    "subparts": ("uda.b200.shuffle.subpartitions", int),
    "push_ctas": ("uda.b200.shuffle.push.ctas", int),
    "merge_ctas_per_sm": ("uda.b200.merge.ctas.per.sm", int),
    "arena_slots": ("uda.b200.shuffle.arena.slots", int),
    "device_cuts": ("uda.b200.shuffle.device.cuts", bool),
    "local_copy": ("uda.b200.shuffle.local.copy", str),
    "verify_crc": ("uda.b200.shuffle.verify.crc", bool),
    "wait_max_spins": ("uda.b200.shuffle.wait.max.spins", int),
}

_CODEC_BLOCK_KEY = {"lzo": "lzo_buffersize", "snappy": "snappy_buffersize"}


def _parse_bool(v) -> bool:
    if isinstance(v, bool):
        return v
    return str(v).strip().lower() in ("1", "true", "yes", "on")


def _parse_xmx(opts: str) -> int:
    """Heap size in bytes from a ``-Xmx`` option, 0 when absent (UdaPlugin.java:216-224 does the same scan)."""
    for tok in opts.split():
        if tok.startswith("-Xmx"):
            v = tok[4:]
            mult = {"k": 1 << 10, "m": 1 << 20, "g": 1 << 30}.get(v[-1:].lower())
            return int(v[:-1]) * mult if mult else int(v)
    return 0


@dataclass
class UdaConfig:
    wqe_per_conn: int = 256
    cma_port: int = 9011
    merge_approach: int = 1                 # 1 online, 2 hybrid, 3 device merge (0 is parsed and rejected like the reference)
    buf_size_kb: int = 1024
    buf_size_min_kb: int = 16
    shuffle_total_size: int = 0             # 0 -> Xmx x shuffle_input_buffer_percent
    shuffle_input_buffer_percent: float = 0.7
    hybrid_lpq_size: int = 0                # 0 -> sqrt(numMaps)
    num_parallel_lpqs: int = 0              # 0 -> 3
    compression_buffer_ratio: float = 0.20
    mem_use_contig_pages: int = 1
    map_output_compression_codec: str = ""
    compress_map_output: bool = False
    lzo_buffersize: int = 256 * 1024
    snappy_buffersize: int = 256 * 1024
    lzo_decompressor: str = "LZO1X"
    log_to_unique_file: bool = False
    developer_mode: bool = False
    blocked_threads_per_disk: int = 1
    provider_reader: str = "aio"            # aio | blocked (csrc/runtime/reader.cc)
    decompress_ahead: bool = True           # compressed map outputs: decode the next block of every segment on the reduce task's decompress thread
    local_dirs: List[str] = field(default_factory=lambda: ["/tmp"])
    map_output_key_class: str = "org.apache.hadoop.io.Text"
    child_java_opts: str = "-Xmx200m"
    hbm_budget_bytes: int = 0               # 0 -> 80 % of the free HBM at job set-up
    subparts: int = 0                       # 0 -> 4 on one GPU, 8 on two, 12 on more
    push_ctas: int = 0                      # 0 -> measured default for the GPU count
    merge_ctas_per_sm: int = 0
    arena_slots: int = 0                    # 0 -> one receive slot per sub-partition; fewer = ring with credits (several GPUs)
    device_cuts: bool = False               # cut the sub-partitions on the device (map outputs with Hadoop's P partitions)
    local_copy: str = "auto"                # auto | push | ce: how a rank's own map outputs reach its arena
    verify_crc: bool = False
    wait_max_spins: int = 50_000_000
    extra: Dict[str, str] = field(default_factory=dict)   # keys this object does not know: kept for getConfData

    # ---------------------------------------------------------------------------------------- construction
    @classmethod
    def from_dict(cls, conf: Mapping[str, object]) -> "UdaConfig":
        by_key = {k: (f, t) for f, (k, t) in _KEYS.items()}
        kw, extra = {}, {}
        for k, v in conf.items():
            if k in by_key:
                f, t = by_key[k]
                if t is bool:
                    kw[f] = _parse_bool(v)
                elif t is list:
                    kw[f] = list(v) if isinstance(v, (list, tuple)) else [x for x in str(v).split(",") if x]
                else:
                    kw[f] = t(v)
            else:
                extra[k] = str(v)
        c = cls(**kw)
        c.extra = extra
        c.validate()
        return c

    @classmethod
    def from_hadoop_xml(cls, *paths: str) -> "UdaConfig":
        """Reads ``<configuration><property><name/><value/></property>…`` files (mapred-site.xml, yarn-site.xml);
        later files win, as in Hadoop's resource order."""
        conf: Dict[str, str] = {}
        for p in paths:
            for prop in ET.parse(p).getroot().iter("property"):
                name, value = prop.findtext("name"), prop.findtext("value")
                if name is not None and value is not None:
                    conf[name.strip()] = value.strip()
        return cls.from_dict(conf)

    def to_dict(self) -> Dict[str, str]:
        out = dict(self.extra)
        for f, (k, t) in _KEYS.items():
            v = getattr(self, f)
            out[k] = ",".join(v) if t is list else ("true" if v else "false") if t is bool else str(v)
        return out

    def validate(self) -> None:
        if self.merge_approach not in (0, 1, 2, 3):
            raise ValueError("mapred.netmerger.merge.approach must be 0, 1 (online), 2 (hybrid) or 3 (device merge)")
        if self.buf_size_kb < self.buf_size_min_kb:
            raise ValueError("mapred.rdma.buf.size is below mapred.rdma.buf.size.min")
        if not (0.0 < self.shuffle_input_buffer_percent <= 1.0):
            raise ValueError("mapred.job.shuffle.input.buffer.percent must be in (0, 1]")
        if self.wqe_per_conn < 2 or not (0 < self.cma_port < 65536):
            raise ValueError("bad mapred.rdma.wqe.per.conn / mapred.rdma.cma.port")
        if self.local_copy not in ("auto", "push", "ce"):
            raise ValueError("uda.b200.shuffle.local.copy must be auto, push or ce")
        if self.arena_slots < 0 or (self.subparts > 0 and self.arena_slots > self.subparts):
            raise ValueError("uda.b200.shuffle.arena.slots must be between 0 and the number of sub-partitions")

    # ---------------------------------------------------------------------------------------- derived values
    def get(self, name: str, default: str = "") -> str:
        """``getConfData(name, default)`` as the C++ side sees it."""
        return self.to_dict().get(name, default)

    @property
    def codec_short(self) -> str:
        if not self.compress_map_output or not self.map_output_compression_codec:
            return ""
        c = self.map_output_compression_codec.lower()
        return "lzo" if "lzo" in c else "snappy" if "snappy" in c else "unknown"

    def codec_block_size(self) -> int:
        k = _CODEC_BLOCK_KEY.get(self.codec_short)
        return getattr(self, k) if k else 0

    def shuffle_memory_bytes(self) -> int:
        """Explicit total, else Xmx x percent (UdaPlugin.java:209-231)."""
        if self.shuffle_total_size > 0:
            return self.shuffle_total_size
        return int(_parse_xmx(self.child_java_opts) * self.shuffle_input_buffer_percent)

    def lpq_size(self, num_maps: int) -> int:
        return self.hybrid_lpq_size if self.hybrid_lpq_size > 0 else max(1, int(math.sqrt(num_maps)))

    def parallel_lpqs(self) -> int:
        return max(3, self.num_parallel_lpqs) if self.num_parallel_lpqs else 3

    # ---------------------------------------------------------------------------------------- wire forms
    def consumer_argv(self, log_dir: str, log_threshold: Optional[int] = None) -> List[str]:
        """argv of the NetMerger role (UdaPluginRT: ``-w -r -a -m 1 -g <dir> -s``)."""
        a = ["-w", str(self.wqe_per_conn), "-r", str(self.cma_port), "-a", str(self.merge_approach), "-m", "1",
             "-g", log_dir, "-s", str(self.buf_size_kb)]
        if log_threshold is not None:
            a += ["-t", str(log_threshold)]
        return a

    def provider_argv(self, log_dir: str, log_threshold: Optional[int] = None) -> List[str]:
        """argv of the MOFSupplier role (UdaShuffleProviderPluginShared.java:29-46: ``-w -r -m 1 -g <dir> -s``)."""
        a = ["-w", str(self.wqe_per_conn), "-r", str(self.cma_port), "-m", "1", "-g", log_dir, "-s", str(self.buf_size_kb)]
        if log_threshold is not None:
            a += ["-t", str(log_threshold)]
        return a

    def init_params(self, num_maps: int, job_id: str, reduce_attempt_id: str, local_dirs: Optional[Sequence[str]] = None) -> List[str]:
        """Parameter list of the INIT command (UdaPlugin.java:269-311 <-> src/Merger/reducer.cc:56-99)."""
        dirs = list(local_dirs if local_dirs is not None else self.local_dirs)
        codec = self.map_output_compression_codec if self.codec_short else "null"
        return [str(num_maps), job_id, reduce_attempt_id, str(self.hybrid_lpq_size), str(self.buf_size_kb * 1024),
                str(self.buf_size_min_kb * 1024), self.map_output_key_class, codec, str(self.codec_block_size()),
                str(self.shuffle_memory_bytes()), str(len(dirs)), *dirs]

    def terasort_job_kwargs(self, world: int) -> Dict[str, object]:
        """Keyword arguments for :class:`uda_b200.models.terasort.TeraSortJob` derived from the ``uda.b200.*`` keys:
        zeros mean "measured default for this many GPUs", the same defaults ``bench.py`` uses (docs/TUNING.md): one GPU
        load/store push with 4 pipelined sub-partitions and the local segments on the copy engines; two GPUs copy-engine
        push on 32 SMs, 8 sub-partitions; more 12 sub-partitions on 36 SMs (48 on 8 GPUs)."""
        multi = world > 1
        subparts = self.subparts if self.subparts > 0 else ((12 if world > 2 else 8) if multi else (1 if self.local_copy == "push" else 4))
        mode = "tma" if multi else "lsu"
        if self.arena_slots and not multi:
            raise ValueError("uda.b200.shuffle.arena.slots needs the copy-engine push, which serves several GPUs")
        if self.arena_slots > subparts:
            raise ValueError("uda.b200.shuffle.arena.slots exceeds the number of sub-partitions")
        if self.push_ctas > 0:
            ctas = self.push_ctas
        elif subparts <= 1:
            ctas = 0
        elif mode == "tma":
            ctas = 32 if world == 2 else (48 if world >= 8 else 36)
        else:
            ctas = 296
        return {"subparts": subparts, "push_mode": mode, "push_ctas": ctas, "merge_ctas_per_sm": self.merge_ctas_per_sm,
                "arena_slots": self.arena_slots, "device_cuts": self.device_cuts, "local_copy": self.local_copy}

    def replace(self, **kw) -> "UdaConfig":
        c = dataclasses.replace(self, **kw)
        c.validate()
        return c
