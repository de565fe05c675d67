"""Compare the SASS of two builds of one object kernel by kernel (addresses and immediates that are
addresses normalised): `python tools/sass_compare.py old.o new.o`.  Used at the end of round 1 to show that
adding the opt-in kernel forms left the measured default kernels unchanged."""
import collections
import hashlib
import re
import subprocess
import sys


def sass(obj):
    out = subprocess.run(['cuobjdump', '-sass', obj], capture_output=True, text=True, check=True).stdout
    funcs = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.search(r'Function : (\S+)', line)
        if m:
            cur = m.group(1)
            funcs[cur] = []
            continue
        m = re.match(r'\s+/\*[0-9a-f]+\*/\s+(.*?);', line)
        if m and cur:
            funcs[cur].append(re.sub(r'0x[0-9a-f]+', 'X', m.group(1).strip()))
    return funcs


def digest(v):
    return hashlib.md5('\n'.join(v).encode()).hexdigest()[:10]


def short(k):
    return re.sub(r'_ZN\d+_GLOBAL__N__[0-9a-f]+_\d+_\w+?_cu_[0-9a-f]{8}\d\d', '', k)[:56]


def main():
    old, new = sass(sys.argv[1]), sass(sys.argv[2])
    oh = {digest(v): k for k, v in old.items()}
    for k, v in new.items():
        print('| `%s` | %d | %s |' % (short(k), len(v), 'identical to `%s`' % short(oh[digest(v)]) if digest(v) in oh else 'new'))


if __name__ == '__main__':
    main()
