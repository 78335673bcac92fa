#!/usr/bin/env python
"""tests/golden/decoder_vectors.npz -> a directory a Rust test can read with std alone:
   cap_NN.i16 (little-endian 16-bit PCM, sample = value / 32768.0 as f32), ask_NN.i16, expected.json.
usage: python tools/export_golden_raw.py <out dir>"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main(out):
    os.makedirs(out, exist_ok=True)
    z = np.load(os.path.join(ROOT, "tests", "golden", "decoder_vectors.npz"))
    exp = {"captures": [], "chirp": []}
    pos = 0
    for i, n in enumerate(z["cap_len"]):
        z["pcm"][pos:pos + int(n)].astype("<i2").tofile(os.path.join(out, f"cap_{i:02d}.i16"))
        pos += int(n)
        recs = []
        for r in z["records"][z["records"]["stream"] == i]:
            pay = z["payload"][int(r["payload_off"]):int(r["payload_off"]) + int(r["len"])] if r["status"] in (0, 4, 5) else []
            recs.append({"status": int(r["status"]), "preamble_start": int(r["preamble_start"]), "frame_start": int(r["frame_start"]),
                         "consumed": int(r["consumed"]), "hdr_valid": int(r["hdr_valid"]), "len": int(r["len"]), "crc": int(r["crc"]),
                         "type": int(r["type"]), "seq": int(r["seq"]), "src": int(r["src"]), "dst": int(r["dst"]),
                         "payload": [int(b) for b in pay]})
        exp["captures"].append({"file": f"cap_{i:02d}.i16", "line_code": "4b5b" if int(z["cap_kind"][i]) == 1 else "manchester",
                                "samples_per_level": 3, "preamble_bytes": 2, "local_addr": 1,
                                "process_samples_call": int(z["cap_chunk"][i]), "records": recs})
    pos = 0
    for i, n in enumerate(z["ask_len"]):
        z["ask_pcm"][pos:pos + int(n)].astype("<i2").tofile(os.path.join(out, f"ask_{i:02d}.i16"))
        pos += int(n)
        recs = [{"start_index": int(r["start_index"]), "sync_power_bits": int(r["sync_power"].view(np.uint32)), "id": int(r["id"]),
                 "accepted": int(r["accepted"]), "bits": [int(b) for b in np.unpackbits(r["bits"])[:108]]}
                for r in z["ask_records"][z["ask_records"]["stream"] == i]]
        exp["chirp"].append({"file": f"ask_{i:02d}.i16", "records": recs})
    json.dump(exp, open(os.path.join(out, "expected.json"), "w"), indent=1)
    print(out, len(exp["captures"]), "captures,", len(exp["chirp"]), "chirp captures")


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "golden_raw")
