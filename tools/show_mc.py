import json, sys
for f in sys.argv[1:]:
    d = json.load(open(f))
    for k, v in d.items():
        print(f.split("/")[-1], k, "call_ms", round(v["ms"], 3), "kernel_ms", round(v["kernel_ms"], 3), "frac(kernel)", round(v["roofline"]["frac"], 3),
              "frac(call)", round(v["pixels"] * v["roofline"]["alg_bytes_per_px"] / (v["ms"] / 1e3) / 1e9 / v["roofline"]["peak"], 3))
        if "kernel_us_by_block_size" in v:
            print("    ", v["kernel_us_by_block_size"])
