"""Stage the reference's own Python simulator under ``oracle/_ref/`` (TEST / BASELINE INFRASTRUCTURE).

The reference is pure Python, so there is nothing to compile: the "build" of ``oracle/_ref`` is a copy of the files on the
path -- ``simulators/ship_in_transit/{ship_model,ship_engine,controllers,LOS_guidance,utils,obstacle}.py`` and
``RLEnv/MSRL_env_ex.py`` -- taken where they lie under /root/reference.  ``oracle/_ref/`` is git-ignored (no reference
source enters the history) but travels to the GPU box with the snapshot like any other built artefact, so that
``bench.py`` can time the reference's OWN simulator on the box's host cores (cpu_baseline / ``--impl reference``).
Run by ``__graft_entry__.build()`` in the build container; a no-op where /root/reference does not exist.
"""
from __future__ import annotations

import os
import shutil

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("SIT_REFERENCE_SRC", "/root/reference")
REF_DST = os.path.join(_HERE, "_ref")
FILES = [
    "simulators/ship_in_transit/ship_model.py",
    "simulators/ship_in_transit/ship_engine.py",
    "simulators/ship_in_transit/controllers.py",
    "simulators/ship_in_transit/LOS_guidance.py",
    "simulators/ship_in_transit/utils.py",
    "simulators/ship_in_transit/obstacle.py",
    "RLEnv/MSRL_env_ex.py",
]


def stage(force: bool = False) -> str | None:
    """Copy the path's reference files into oracle/_ref/ (same relative layout).  Returns the staged root or None."""
    if not os.path.isdir(os.path.join(REF_SRC, "simulators", "ship_in_transit")):
        return REF_DST if staged() else None
    for rel in FILES:
        src, dst = os.path.join(REF_SRC, rel), os.path.join(REF_DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        if force or not os.path.exists(dst) or os.path.getmtime(dst) < os.path.getmtime(src):
            shutil.copyfile(src, dst)
    with open(os.path.join(REF_DST, "STAGED_FROM"), "w") as f:
        f.write(REF_SRC + "\n" + "\n".join(FILES) + "\n")
    return REF_DST


def staged() -> bool:
    return all(os.path.exists(os.path.join(REF_DST, rel)) for rel in FILES)


if __name__ == "__main__":
    print(stage(force=True))
