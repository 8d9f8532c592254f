"""Canonical op schemas of the REFERENCE, from the m.def(...) string literals of /root/reference/hopper/flash_api.cpp:1672-1761.
Run in the build container only:  python tests/golden/make_golden_schemas.py
Each literal sequence is joined, registered in a scratch torch.library namespace and printed back by torch, so that what is
stored (tests/golden/op_schemas.json) is torch's canonical spelling of the reference's schema; tests/test_abi_cpu.py compares
the whole strings with the ops this repo registers."""
import json
import os
import re

import torch

SRC = "/root/reference/hopper/flash_api.cpp"
HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    text = open(SRC).read()
    body = text[text.index("TORCH_LIBRARY(flash_attn_3, m)"):text.index("TORCH_LIBRARY_IMPL(flash_attn_3, CUDA, m)")]
    lib = torch.library.Library("fa3_reference_schema_scratch", "DEF")
    out = {}
    for m in re.finditer(r"m\.def\(((?:\s*\"[^\"]*\")+)\s*\)", body):
        schema = "".join(re.findall(r"\"([^\"]*)\"", m.group(1)))
        name = schema[:schema.index("(")]
        lib.define(schema)
        canon = str(getattr(torch.ops.fa3_reference_schema_scratch, name).default._schema)
        out[name] = canon.replace("fa3_reference_schema_scratch::", "flash_attn_3::")
    assert set(out) == {"fwd", "bwd", "fwd_combine", "get_scheduler_metadata"}, out.keys()
    json.dump(out, open(os.path.join(HERE, "op_schemas.json"), "w"), indent=1)
    print("wrote", list(out))


if __name__ == "__main__":
    main()


# The FA2-style public signatures (names, order, defaults) of /root/reference/flash_attn/flash_attn_interface.py are stored
# the same way (ast, because that module imports its CUDA extension at import time) in tests/golden/fa2_signatures.json:
#   python - <<'PY'
#   import ast, json
#   out = {}
#   for n in ast.parse(open('/root/reference/flash_attn/flash_attn_interface.py').read()).body:
#       if isinstance(n, ast.FunctionDef) and n.name.startswith('flash_attn'):
#           a = n.args; names = [x.arg for x in a.args]; d = [ast.unparse(x) for x in a.defaults]
#           out[n.name] = [[x, y] for x, y in zip(names, [None] * (len(names) - len(d)) + d)]
#   json.dump(out, open('tests/golden/fa2_signatures.json', 'w'), indent=1)
#   PY
