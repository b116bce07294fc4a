#!/usr/bin/env python
"""Generates tests/golden/jtag_golden.json from the reference's own JTAG driver script,
/root/reference/generators/jtag2mm/src/test/python/gpio_for_jtag.py: its functions jtagInit / jtagSend are executed as they are (the
pyftdi import and the GPIO device at module level are left out: there is no FTDI adapter here) on (a) the script's own command list --
the four register writes it sends to the NCO block, with the intended address / data pairs taken from its arguments -- and (b) a second
list with other values.  The JSON holds the GPIO byte sequences and the bus writes they stand for.  Run in the build container only
(/root/reference does not exist on the GPU box):  python tests/golden/make_jtag_golden.py"""
import ast
import json
import os

SRC = "/root/reference/generators/jtag2mm/src/test/python/gpio_for_jtag.py"
tree = ast.parse(open(SRC).read())
funcs = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("jtagInit", "jtagSend")]
ns = {}
exec(compile(ast.Module(body=funcs, type_ignores=[]), SRC, "exec"), ns)
jtagInit, jtagSend = ns["jtagInit"], ns["jtagSend"]

# the module-level command list of the script (lines 58-84), replayed call by call
calls = [n.value for n in tree.body if isinstance(n, ast.Assign) and isinstance(n.value, ast.Call)
         and getattr(n.value.func, "id", "") in ("jtagInit", "jtagSend")]


def run(call_list):
    sig = []
    for c in call_list:
        args = [ast.literal_eval(a) for a in c.args[1:]]
        sig = ns[c.func.id](sig, *args)
    return sig


def writes_of(call_list):
    """(address, data) of every `0x03 data / 0x02 address / 0x01 write` triple, from the arguments of the calls"""
    out, data, addr = [], None, None
    pending = None
    for c in call_list:
        if c.func.id != "jtagSend":
            continue
        val, length, is_data = [ast.literal_eval(a) for a in c.args[1:]]
        if not is_data:
            pending = val
            if val == 0x01:
                out.append([addr, data])
        elif pending == 0x03:
            data = val
        elif pending == 0x02:
            addr = val
    return out


script = run(calls)
cases = [{"name": "gpio_for_jtag.py command list", "ir_length": 3, "samples": script, "writes": writes_of(calls)}]

# (b) the same functions on other values: a 32-bit pattern with both end bits set, the highest 16-bit address
sig = jtagInit([])
want = []
for data, addr in ((0x80000001, 0xFFFC), (0x12345678, 0x0100), (0xDEADBEEF, 0x2008)):
    sig = jtagSend(sig, 0x03, 3, False); sig = jtagSend(sig, data, 32, True)
    sig = jtagSend(sig, 0x02, 3, False); sig = jtagSend(sig, addr, 16, True)
    sig = jtagSend(sig, 0x01, 3, False)
    want.append([addr, data])
cases.append({"name": "jtagSend on other values", "ir_length": 3, "samples": sig, "writes": want})

out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "jtag_golden.json")
json.dump({"source": SRC, "pins": {"tck": 0x10, "tms": 0x20, "tdi": 0x40}, "cases": cases}, open(out, "w"))
print(out, [len(c["samples"]) for c in cases], cases[0]["writes"])
