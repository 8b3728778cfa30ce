"""Documentation that is generated from code must match the code: the flag reference (docs/parameter.md) is the
output of tools/gen_param_docs.py over the real parsers, and the C header declares every entry point that the ctypes
binding expects (ops/signatures.py), so neither can go stale unnoticed. The reference's flag documentation is a
hand-written Sphinx page (/root/reference/docs/parameter.rst)."""
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_parameter_reference_is_current():
    r = subprocess.run([sys.executable, os.path.join(ROOT, 'tools', 'gen_param_docs.py')], capture_output=True, text=True,
                       timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    have = open(os.path.join(ROOT, 'docs', 'parameter.md')).read()
    assert r.stdout.strip() == have.strip(), 'docs/parameter.md is stale: python tools/gen_param_docs.py > docs/parameter.md'


def test_header_declares_every_bound_entry_point():
    from hetseq_b200.ops import signatures
    header = open(os.path.join(ROOT, 'include', 'hetseq_b200.h')).read()
    declared = set(re.findall(r'\b(hs_[a-z0-9_]+)\s*\(', header))
    missing = sorted(set(signatures.SIGNATURES) - declared)
    assert not missing, 'bound in ops/signatures.py but not declared in include/hetseq_b200.h: {}'.format(missing)
    # argument counts agree as well (a changed signature must be changed in both places)
    flat = re.sub(r'/\*.*?\*/', ' ', header, flags=re.S)
    for name, argtypes in signatures.SIGNATURES.items():
        m = re.search(r'\b%s\s*\(([^;]*?)\)\s*;' % re.escape(name), flat, flags=re.S)
        assert m, name
        args = m.group(1).strip()
        n = 0 if args in ('', 'void') else len([a for a in args.split(',') if a.strip()])
        assert n == len(argtypes), '{}: header has {} parameters, ops/signatures.py {}'.format(name, n, len(argtypes))


def test_library_exports_exactly_the_declared_abi():
    """The built library exports every function the header declares and nothing else under the hs_ prefix."""
    import shutil
    import pytest
    from hetseq_b200.ops.build import CUDA_LIB
    if not CUDA_LIB.exists() or shutil.which('nm') is None:
        pytest.skip('library not built (python -c "import __graft_entry__ as g; g.build()") or no nm')
    out = subprocess.run(['nm', '-D', '--defined-only', str(CUDA_LIB)], capture_output=True, text=True, check=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if l.split() and l.split()[-1].startswith('hs_')}
    header = open(os.path.join(ROOT, 'include', 'hetseq_b200.h')).read()
    declared = set(re.findall(r'\b(hs_[a-z0-9_]+)\s*\(', header))
    assert exported == declared, (sorted(exported - declared), sorted(declared - exported))
