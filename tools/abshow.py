"""Prints the step time and the top kernels of bench.py JSON lines: python tools/abshow.py FILE [FILE ...]"""
import json, sys
for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    k = d.get("kernels", {})
    names = ["k_blend_bwd", "k_deform_bwd_weights", "k_deform_fwd_tc", "k_deform_bwd_data_saved", "k_blend_fwd", "k_deform_sample_bwd",
             "k_tile_sort", "k_preprocess_bwd", "k_scatter", "k_preprocess"]
    print(f.split("/")[-1], d["ms_per_step"], "ms", d["value"], "it/s e2e", d["e2e"]["value"], "|",
          " ".join(f"{n[2:]}={k[n]['avg_us']:.0f}" for n in names if n in k))
