#!/usr/bin/env python3
"""A stand-in for ``hadoop fs`` used by the tests: the same sub-commands and output shapes over a local
directory (``$FAKE_HDFS_ROOT`` + the path with its ``hdfs:`` / ``afs:`` scheme stripped)."""
import os
import shutil
import sys
import time


def local(p):
    for scheme in ("hdfs:", "afs:"):
        if p.startswith(scheme):
            p = p[len(scheme):]
    while p.startswith("//"):
        p = p[1:]
    return os.path.join(os.environ["FAKE_HDFS_ROOT"], p.lstrip("/"))


def main(argv):
    assert argv and argv[0] == "fs", argv
    args = argv[1:]
    log = os.environ.get("FAKE_HDFS_LOG")
    if log:
        with open(log, "a") as f:
            f.write(" ".join(argv) + "\n")
    while args and (args[0].startswith("-D") or args[0] == "-conf"):
        args = args[2:] if args[0] == "-conf" else args[1:]
    cmd, rest = args[0], args[1:]
    if cmd == "-test":
        flag, p = rest[0], local(rest[1])
        ok = {"-e": os.path.exists, "-d": os.path.isdir, "-f": os.path.isfile}[flag](p)
        return 0 if ok else 1
    if cmd == "-ls":
        p = local(rest[0])
        if not os.path.exists(p):
            print(f"ls: `{rest[0]}': No such file or directory", file=sys.stderr)
            return 1
        names = sorted(os.listdir(p)) if os.path.isdir(p) else [""]
        print(f"Found {len(names)} items")
        for n in names:
            q = os.path.join(p, n) if n else p
            shown = rest[0].rstrip("/") + ("/" + n if n else "")
            perm = "drwxr-xr-x" if os.path.isdir(q) else "-rw-r--r--"
            size = 0 if os.path.isdir(q) else os.path.getsize(q)
            print(f"{perm}   3 user group {size} {time.strftime('%Y-%m-%d %H:%M')} {shown}")
        return 0
    if cmd == "-cat":
        p = local(rest[0])
        if not os.path.isfile(p):
            return 1
        sys.stdout.write(open(p).read())
        return 0
    if cmd == "-mkdir":
        os.makedirs(local(rest[-1]), exist_ok=True)
        return 0
    if cmd in ("-rm", "-rmr"):
        p = local(rest[-1])
        if not os.path.exists(p):
            return 1
        shutil.rmtree(p) if os.path.isdir(p) else os.remove(p)
        return 0
    if cmd == "-touchz":
        p = local(rest[0])
        os.makedirs(os.path.dirname(p), exist_ok=True)
        open(p, "a").close()
        return 0
    if cmd == "-mv":
        os.rename(local(rest[0]), local(rest[1]))
        return 0
    if cmd == "-put":
        rest = [r for r in rest if r != "-f"]
        src, dst = rest[0], local(rest[1])
        if os.path.isdir(dst):
            dst = os.path.join(dst, os.path.basename(src))
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copytree(src, dst) if os.path.isdir(src) else shutil.copy(src, dst)
        return 0
    if cmd == "-get":
        src, dst = local(rest[0]), rest[1]
        if not os.path.exists(src):
            return 1
        shutil.copytree(src, dst) if os.path.isdir(src) else shutil.copy(src, dst)
        return 0
    if cmd == "-du":
        p = local(rest[-1])
        total = sum(os.path.getsize(os.path.join(d, f)) for d, _, fs in os.walk(p) for f in fs) if os.path.isdir(p) \
            else os.path.getsize(p)
        print(f"{total}\t{rest[-1]}")
        return 0
    print("unknown command " + cmd, file=sys.stderr)
    return 2


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
