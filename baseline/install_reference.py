"""Install the unmodified reference into baseline/_ref (git-ignored, shipped to the GPU box).

Order: the prescribed offline pip install; the same with --no-deps; and, because the reference's
build backend (poetry-core) is not in this image or its wheelhouse so neither can build, a
verbatim copy of the reference's ``ddl`` package.  Whatever happened is written to
baseline/_ref/INSTALL_LOG.txt and summarised in DESIGN.md.  No reference file is edited.
"""
import filecmp
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("BDDL_REFERENCE_SRC", "/root/reference")
TARGET = os.path.join(HERE, "_ref")


def main() -> int:
    if not os.path.isdir(REF_SRC):
        print(f"reference source {REF_SRC} not found", file=sys.stderr)
        return 1
    os.makedirs(TARGET, exist_ok=True)
    log = []
    tmp = tempfile.mkdtemp(prefix="refcopy_")
    src = os.path.join(tmp, "reference")
    shutil.copytree(REF_SRC, src, ignore=shutil.ignore_patterns(".git"))
    base = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links",
            "/opt/wheelhouse", "--target", TARGET]
    ok = False
    for extra in ([], ["--no-deps"]):
        res = subprocess.run(base + extra + [src], capture_output=True, text=True)
        tail = (res.stdout + res.stderr).strip().splitlines()[-1:] or [""]
        log.append(f"pip {' '.join(extra) or '(plain)'} -> rc={res.returncode}: {tail[0]}")
        if res.returncode == 0:
            ok = True
            break
    if not ok:
        dst = os.path.join(TARGET, "ddl")
        if os.path.isdir(dst):
            shutil.rmtree(dst)
        shutil.copytree(os.path.join(REF_SRC, "ddl"), dst)
        same = not filecmp.dircmp(os.path.join(REF_SRC, "ddl"), dst).diff_files
        log.append(f"fallback: verbatim copy of {REF_SRC}/ddl -> baseline/_ref/ddl (identical={same})")
    shutil.rmtree(tmp, ignore_errors=True)
    with open(os.path.join(TARGET, "INSTALL_LOG.txt"), "w") as fh:
        fh.write("\n".join(log) + "\n")
    print("\n".join(log))
    return 0


if __name__ == "__main__":
    sys.exit(main())
