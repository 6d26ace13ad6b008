#!/usr/bin/env python3
"""tools/sass_diff.py before.sass after.sass -- which kernels' SASS changed between two builds.

Each input is `cuobjdump -sass nn_bvh_b200/libnnbvh.so`.  Functions are compared instruction for instruction
(encodings and addresses stripped).  Used to show that adding a kernel (e.g. csrc/alpha.cuh) leaves every
measured kernel exactly as it was, so earlier profiles stay valid.
"""
import hashlib
import re
import sys


def functions(path):
    out, name, buf = {}, None, []
    for ln in open(path):
        m = re.search(r"Function : (\S+)", ln)
        if m:
            if name:
                out[name] = hashlib.md5("".join(buf).encode()).hexdigest()
            name, buf = m.group(1), []
        elif name:
            buf.append(re.sub(r"/\*[0-9a-fx]+\*/", "", ln))
    if name:
        out[name] = hashlib.md5("".join(buf).encode()).hexdigest()
    return out


def main():
    a, b = functions(sys.argv[1]), functions(sys.argv[2])
    print(f"{len(a)} functions before, {len(b)} after")
    for title, names in (("changed", [k for k in a if k in b and a[k] != b[k]]),
                         ("new", [k for k in b if k not in a]), ("gone", [k for k in a if k not in b])):
        print(f"{title}: {len(names)}")
        for k in names:
            print("   ", k)


if __name__ == "__main__":
    main()
