"""Compare the SASS of two builds of libmrgpu.so kernel by kernel (instruction text only; addresses and encodings are
ignored): a refactoring that moves device code between headers must leave every kernel unchanged, and this shows it
without a GPU.     cuobjdump -sass old.so > a.sass; cuobjdump -sass new.so > b.sass; python tools/sass_diff.py a.sass b.sass"""
import re
import sys


def load(path):
    fns, cur = {}, None
    for line in open(path):
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            fns[cur] = []
            continue
        if cur is not None:
            m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(.*?)\s*;\s*/\*", line)
            if m:
                fns[cur].append(m.group(1))
    return fns


a, b = load(sys.argv[1]), load(sys.argv[2])
changed = [k for k in a if k in b and a[k] != b[k]]
print(f"{len(a)} / {len(b)} kernels, {len(changed)} changed, {len([k for k in a if k not in b])} gone, {len([k for k in b if k not in a])} new")
for k in changed:
    print("  changed:", k[:140], len(a[k]), "->", len(b[k]), "instructions")
sys.exit(1 if changed or set(a) != set(b) else 0)
