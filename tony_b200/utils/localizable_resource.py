"""Resource strings ``SOURCE[::NAME_IN_CONTAINER][#archive]`` and their localisation.

Grammar from tony-core/src/main/java/com/linkedin/tony/LocalizableResource.java:20-113
(constants Constants.java:79-80).  On YARN "localise" means HDFS download by the NodeManager;
on the local cluster it is a symlink (plain files), a directory copy-by-symlink, or an unzip
into a directory named after the resource (archives) inside the container's working dir
(behaviour checked by the reference's check_archive_file_localization.py:22-30).
"""
from __future__ import annotations

import os
import shutil
from dataclasses import dataclass
from typing import List, Optional

from ..conf import keys as K
from . import utils as U


class ResourceParseError(ValueError):
    pass


@dataclass
class LocalizableResource:
    complete: str
    source: str
    local_name: str
    is_archive: bool
    is_directory: bool

    @staticmethod
    def parse(resource: str, must_exist: bool = True) -> "LocalizableResource":
        complete = resource.strip()
        path = complete
        archive = False
        if complete.lower().endswith(K.ARCHIVE_SUFFIX):
            archive = True
            path = complete[: -len(K.ARCHIVE_SUFFIX)]
        parts = path.split(K.RESOURCE_DIVIDER)
        if len(parts) > 2 or not parts[0]:
            raise ResourceParseError(f"Failed to parse file: {complete}")
        source = parts[0]
        if source.startswith("file://"):
            source = source[len("file://"):]
        if must_exist and not os.path.exists(source):
            raise FileNotFoundError(source)
        local_name = parts[1] if len(parts) == 2 else os.path.basename(source.rstrip("/"))
        return LocalizableResource(complete, source, local_name, archive, os.path.isdir(source))

    def to_string(self, new_source: Optional[str] = None) -> str:
        s = (new_source or self.source) + K.RESOURCE_DIVIDER + self.local_name
        return s + K.ARCHIVE_SUFFIX if self.is_archive else s

    def localize(self, container_dir: str) -> str:
        """Materialise this resource inside ``container_dir``; returns the local path."""
        dst = os.path.join(container_dir, self.local_name)
        if os.path.lexists(dst):
            return dst
        if self.is_archive:
            U.unzip_archive(self.source, dst)
        elif self.is_directory:
            # a directory contributes its first-level files (TonyClient.java:726-770)
            os.makedirs(dst, exist_ok=True)
            for name in os.listdir(self.source):
                full = os.path.join(self.source, name)
                if os.path.isfile(full):
                    os.symlink(os.path.abspath(full), os.path.join(dst, name))
        else:
            try:
                os.symlink(os.path.abspath(self.source), dst)
            except OSError:
                shutil.copy2(self.source, dst)
        return dst


def parse_resources(csv: Optional[str]) -> List[LocalizableResource]:
    out = []
    for item in (csv or "").split(","):
        if item.strip():
            out.append(LocalizableResource.parse(item))
    return out


def stage_resource(resource: str, staging_dir: str) -> str:
    """Client side: copy/zip a resource into the application's staging directory and return
    the rewritten resource string (local directories become archives, like the reference)."""
    r = LocalizableResource.parse(resource)
    os.makedirs(staging_dir, exist_ok=True)
    if r.is_directory:
        zip_path = os.path.join(staging_dir, r.local_name + ".zip")
        U.zip_folder(r.source, zip_path)
        return LocalizableResource(r.complete, zip_path, r.local_name, True, False).to_string()
    dst = os.path.join(staging_dir, os.path.basename(r.source))
    if os.path.abspath(dst) != os.path.abspath(r.source):
        shutil.copy2(r.source, dst)
    return LocalizableResource(r.complete, dst, r.local_name, r.is_archive, False).to_string()
