"""CPU-side checks of the drop-in boundary: libstage_gpu.so loads and exports exactly the symbols
include/stage_gpu.h declares (no compute calls: there is no GPU here)."""
import ctypes
import os
import re

import stage_indexorganized_b200 as sg

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "stage_gpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(stage_gpu_\w+)\s*\(", src)))


def test_library_is_built_in_tree():
    assert os.path.exists(sg.LIB_PATH), "run __graft_entry__.build()"


def test_every_declared_symbol_is_exported():
    lib = ctypes.CDLL(sg.LIB_PATH)
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/stage_gpu.h but not exported"


def test_python_binding_covers_the_header():
    assert sorted(sg.SYMBOLS) == _declared()


def test_strerror_and_route_host_only():
    import numpy as np
    l = sg.lib()
    assert l.stage_gpu_strerror(0) == b"ok"
    assert b"quiesced" in l.stage_gpu_strerror(-3)
    # routing is host-only logic: YCSB int keys order is signed-byte, LSB first (b_tree.h:99-134)
    keys = np.array([0, 1, 255, 256, 128, 32896], np.uint32).view(np.uint8).reshape(-1, 4)
    split = np.array([0], np.uint32).view(np.uint8).reshape(-1, 4)       # shard 0 owns keys <= 0
    owner = sg.route(keys, split, 4)
    # in reference order 128 (0x80 -> -128) and 32896 sort before 0; 1, 255(-1)?, 256 ...
    #   byte0 of 255 is 0xFF = -1 < 0  -> before key 0 ; byte0 of 1 = 1 > 0 -> after
    assert owner.tolist() == [0, 1, 0, 1, 0, 0]


def test_header_is_plain_c_and_links(tmp_path):
    """include/stage_gpu.h is the boundary a cgo / JNI / ctypes binding would read: it must compile as C99 (no C++ in
    the signatures) and a C program must link against the library and call the host-only entry points."""
    import subprocess
    src = tmp_path / "abi.c"
    src.write_text(r'''
#include <stdio.h>
#include <string.h>
#include "stage_gpu.h"
int main(void) {
    stage_schema_t sc = {8, 8, 4096, STAGE_KEY_VARLEN};
    unsigned char slot[8], key[8];
    uint32_t len = 0;
    if (strcmp(stage_gpu_strerror(STAGE_OK), "ok") != 0) return 1;
    if (stage_gpu_key_pack(&sc, "4711", 4, slot) != STAGE_OK) return 2;
    if (stage_gpu_key_unpack(&sc, slot, key, &len) != STAGE_OK || len != 4 || memcmp(key, "4711", 4) != 0) return 3;
    if (stage_gpu_key_pack(&sc, "12345678", 8, slot) != STAGE_ERR_ARG) return 4;      /* longer than slot - 1 */
    if (stage_gpu_set_row_slack(-1.0) != STAGE_ERR_ARG || stage_gpu_set_row_slack(0.0) != STAGE_OK) return 5;
    printf("c-abi ok\n");
    return 0;
}
''')
    exe = tmp_path / "abi"
    libdir = os.path.dirname(sg.LIB_PATH)
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"), str(src),
                        "-o", str(exe), "-L", libdir, "-lstage_gpu", f"-Wl,-rpath,{libdir}"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0 and "c-abi ok" in r.stdout, (r.returncode, r.stdout, r.stderr)
