"""Hadoop-style XML configuration: reader, writer and the layered TonY configuration.

Precedence (lowest to highest), as in the reference client
(tony-core/src/main/java/com/linkedin/tony/TonyClient.java:657-691):
``tony-default.xml`` -> ``--conf_file`` (else ``./tony.xml``) -> repeated ``--conf k=v``
(multi-valued keys append) -> ``$TONY_CONF_DIR/tony-site.xml``.  The merged result is
frozen into ``tony-final.xml`` that the ApplicationMaster and every TaskExecutor re-read
(ApplicationMaster.java:230, TaskExecutor.java:257) and that the history portal parses
(util/ParserUtils.java:181-240).
"""
from __future__ import annotations

import os
import re
import xml.etree.ElementTree as ET
from pathlib import Path
from typing import Dict, Iterable, Iterator, List, Optional, Tuple
from xml.sax.saxutils import escape

from . import keys as K

DEFAULT_XML_PATH = Path(__file__).resolve().parent / K.TONY_DEFAULT_XML
_MEM_RE = re.compile(r"^\s*(\d+)\s*([mMgG]?)\s*$")
_TRUE = {"true", "1", "yes", "on"}


def parse_memory_string(mem: str) -> int:
    """'2g' -> 2048, '512m' -> 512, '1024' -> 1024 (MB).  Reference: util/Utils.java:152-163."""
    m = _MEM_RE.match(str(mem))
    if not m:
        raise ValueError(f"Failed to parse memory string: {mem!r}")
    n = int(m.group(1))
    return n * 1024 if m.group(2).lower() == "g" else n


def read_xml(path) -> List[Tuple[str, str, Dict[str, str]]]:
    """Return [(name, value, attrs)] from a Hadoop configuration XML file."""
    root = ET.parse(str(path)).getroot()
    out = []
    for prop in root.iter("property"):
        name = (prop.findtext("name") or "").strip()
        if not name:
            continue
        value = prop.findtext("value")
        attrs = {}
        for tag in ("final", "source", "description"):
            t = prop.findtext(tag)
            if t is not None:
                attrs[tag] = t.strip()
        out.append((name, (value or "").strip(), attrs))
    return out


def write_xml(path, items: Iterable[Tuple[str, str, Dict[str, str]]]) -> None:
    lines = ['<?xml version="1.0" encoding="UTF-8" standalone="no"?>', "<configuration>"]
    for name, value, attrs in items:
        lines.append("  <property>")
        lines.append(f"    <name>{escape(name)}</name>")
        lines.append(f"    <value>{escape(value)}</value>")
        for tag in ("final", "source", "description"):
            if attrs and tag in attrs:
                lines.append(f"    <{tag}>{escape(attrs[tag])}</{tag}>")
        lines.append("  </property>")
    lines.append("</configuration>")
    tmp = str(path) + ".tmp"
    with open(tmp, "w", encoding="utf-8") as f:
        f.write("\n".join(lines) + "\n")
    os.replace(tmp, str(path))


class TonyConfiguration:
    """Ordered key -> string map with typed getters and source tracking."""

    def __init__(self, load_defaults: bool = True):
        self._values: Dict[str, str] = {}
        self._source: Dict[str, str] = {}
        if load_defaults:
            self.add_resource(DEFAULT_XML_PATH)

    # -- loading -----------------------------------------------------------------------
    def add_resource(self, path) -> "TonyConfiguration":
        src = os.path.basename(str(path))
        for name, value, _ in read_xml(path):
            self._values[name] = value
            self._source[name] = src
        return self

    def set(self, key: str, value, source: str = "programmatically") -> None:
        self._values[key] = str(value)
        self._source[key] = source

    def set_cli(self, key: str, value: str) -> None:
        """``--conf k=v``: multi-valued keys append, everything else overrides."""
        if key in K.MULTI_VALUE_KEYS and self._values.get(key):
            self.set(key, self._values[key] + "," + value, "cli")
        else:
            self.set(key, value, "cli")

    def unset(self, key: str) -> None:
        self._values.pop(key, None)
        self._source.pop(key, None)

    # -- typed access ------------------------------------------------------------------
    def get(self, key: str, default: Optional[str] = None) -> Optional[str]:
        v = self._values.get(key)
        return default if v is None else v

    def get_trimmed(self, key: str, default: Optional[str] = None) -> Optional[str]:
        v = self.get(key, default)
        return v.strip() if isinstance(v, str) else v

    def get_int(self, key: str, default: int = 0) -> int:
        v = self.get(key)
        if v is None or v.strip() == "":
            return default
        return int(v.strip())

    def get_float(self, key: str, default: float = 0.0) -> float:
        v = self.get(key)
        if v is None or v.strip() == "":
            return default
        return float(v.strip())

    def get_bool(self, key: str, default: bool = False) -> bool:
        v = self.get(key)
        if v is None or v.strip() == "":
            return default
        return v.strip().lower() in _TRUE

    def get_strings(self, key: str, default: Optional[List[str]] = None) -> List[str]:
        v = self.get(key)
        if v is None or v.strip() == "":
            return list(default or [])
        return [s.strip() for s in v.split(",") if s.strip()]

    def get_memory_mb(self, key: str, default: str = K.DEFAULT_MEMORY) -> int:
        return parse_memory_string(self.get(key, default) or default)

    def __contains__(self, key: str) -> bool:
        return key in self._values

    def __iter__(self) -> Iterator[Tuple[str, str]]:
        return iter(self._values.items())

    def items(self):
        return self._values.items()

    def source_of(self, key: str) -> Optional[str]:
        return self._source.get(key)

    def copy(self) -> "TonyConfiguration":
        c = TonyConfiguration(load_defaults=False)
        c._values = dict(self._values)
        c._source = dict(self._source)
        return c

    # -- job types ----------------------------------------------------------------------
    def job_types(self) -> List[str]:
        """Every ``<job>`` for which ``tony.<job>.instances`` is set, in file order
        (reference: util/Utils.java:451-475 with the regex of TonyConfigurationKeys.java:189)."""
        out = []
        for key in self._values:
            m = K.INSTANCES_REGEX.match(key)
            if m and m.group(1) not in out:
                out.append(m.group(1))
        return out

    def num_instances(self, job: str) -> int:
        return self.get_int(K.instances_key(job), 0)

    def total_instances(self) -> int:
        return sum(self.num_instances(j) for j in self.job_types())

    def untracked_job_types(self) -> List[str]:
        return self.get_strings(K.UNTRACKED_JOBTYPES)

    def sidecar_job_types(self) -> List[str]:
        return self.get_strings(K.SIDECAR_JOBTYPES)

    def is_job_type_monitored(self, job: str) -> bool:
        """Tracked = neither untracked nor sidecar (util/Utils.java:655-682)."""
        return job not in self.untracked_job_types() and job not in self.sidecar_job_types()

    # -- persistence -------------------------------------------------------------------
    def write_xml(self, path) -> None:
        write_xml(path, ((k, v, {"source": self._source.get(k, "")}) for k, v in self._values.items()))

    @staticmethod
    def from_final_xml(path) -> "TonyConfiguration":
        c = TonyConfiguration(load_defaults=False)
        for name, value, attrs in read_xml(path):
            c._values[name] = value
            c._source[name] = attrs.get("source", os.path.basename(str(path)))
        return c


def build_configuration(conf_file: Optional[str] = None, cli_confs: Optional[Iterable[str]] = None,
                        cwd: Optional[str] = None, env: Optional[Dict[str, str]] = None) -> TonyConfiguration:
    """Layer the four sources in reference order."""
    env = os.environ if env is None else env
    conf = TonyConfiguration()
    if conf_file:
        conf.add_resource(conf_file)
    else:
        local = Path(cwd or os.getcwd()) / K.TONY_XML
        if local.exists():
            conf.add_resource(local)
    for kv in cli_confs or []:
        if "=" not in kv:
            raise ValueError(f"--conf expects key=value, got {kv!r}")
        k, v = kv.split("=", 1)
        conf.set_cli(k.strip(), v.strip())
    conf_dir = env.get(K.TONY_CONF_DIR_ENV)
    if conf_dir:
        site = Path(conf_dir) / K.TONY_SITE_XML
        if site.exists():
            conf.add_resource(site)
    return conf


def write_default_xml(path=DEFAULT_XML_PATH) -> None:
    """(Re)generate tony-default.xml from keys.DEFAULTS — the single source of truth."""
    write_xml(path, ((k, v, {"description": d}) for k, v, d in K.DEFAULTS))
