"""File/directory permission probe (parity: /root/reference/pytorch_impl/libs/tools/access.py:25-62)."""

from __future__ import annotations

import os
import pathlib
import stat

__all__ = ["can_access"]


def _mode_allows(st, read, write):
    if os.geteuid() == st.st_uid:
        rbit, wbit = stat.S_IRUSR, stat.S_IWUSR
    elif os.getegid() == st.st_gid:
        rbit, wbit = stat.S_IRGRP, stat.S_IWGRP
    else:
        rbit, wbit = stat.S_IROTH, stat.S_IWOTH
    return (not read or bool(st.st_mode & rbit)) and (not write or bool(st.st_mode & wbit))


def can_access(path, read=False, write=False, recurse=False):
    """True iff `path` exists and the file (or every file below the directory; sub-directories
    only when `recurse`) grants the requested permission bits to the effective user."""
    try:
        path = pathlib.Path(path)
        if not path.exists():
            return False
        st = path.stat()
        if not stat.S_ISDIR(st.st_mode):
            return _mode_allows(st, read, write)
        for child in path.iterdir():
            if child.is_dir() and not recurse:
                continue
            if not can_access(child, read, write, recurse):
                return False
        return True
    except OSError:
        return False
