"""Compare the SASS of every kernel in two builds of libps_engine.so (or any two cubins / .so files).

    python tools/sass_diff.py OLD.so NEW.so [> profiles/rNN_sass_diff_<what>.md]

Used when a change must not touch kernels that were already verified on hardware and no GPU is at hand: identical
instruction streams (opcodes + operands, addresses and encodings stripped) mean the verified results carry over.
"""
import collections
import hashlib
import re
import subprocess
import sys


def kernels(path):
    out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True, check=True).stdout
    fn, body, res = None, [], collections.OrderedDict()
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            if fn:
                res[fn] = body
            fn, body = m.group(1), []
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(.*?)\s*/\* 0x[0-9a-f]+ \*/", line)
        if m and fn:
            body.append(m.group(1).strip())
    if fn:
        res[fn] = body
    return res


def demangle(names):
    try:
        out = subprocess.run(["cu++filt"] + list(names), capture_output=True, text=True, check=True).stdout.splitlines()
        return dict(zip(names, out))
    except Exception:
        return {n: n for n in names}


def short(s, n=150):
    s = re.sub(r"ps::Traits<[^>]*>", "Traits<..>", s)
    return s if len(s) <= n else s[:n] + "..."


def main():
    a, b = kernels(sys.argv[1]), kernels(sys.argv[2])
    names = demangle(sorted(set(a) | set(b)))
    print(f"| kernel | instructions old | new | same stream |\n|---|---|---|---|")
    same = diff = 0
    for k in sorted(set(a) | set(b), key=lambda x: names[x]):
        ia, ib = a.get(k), b.get(k)
        if ia is None or ib is None:
            print(f"| `{short(names[k])}` | {len(ia) if ia else '-'} | {len(ib) if ib else '-'} | only in {'old' if ib is None else 'new'} |")
            continue
        eq = ia == ib
        same += eq; diff += not eq
        tag = "yes" if eq else f"NO (sha {hashlib.sha1(chr(10).join(ia).encode()).hexdigest()[:8]} vs {hashlib.sha1(chr(10).join(ib).encode()).hexdigest()[:8]})"
        print(f"| `{short(names[k])}` | {len(ia)} | {len(ib)} | {tag} |")
    print(f"\n{same} kernels identical, {diff} differ.")


if __name__ == "__main__":
    main()
