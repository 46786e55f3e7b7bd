#!/usr/bin/env python
"""sweep_table.py <bench line json> [<N=2 bench line json>]: the markdown tables of DESIGN.md section 6.1, generated from what
bench.py printed (headline, roofline, e2e, cpu_baseline, sweep, pingpong, himeno) -- nothing in them is typed by hand.
  python tools/sweep_table.py profiles/bench/r2_bench_n1_final.json profiles/bench/r2_bench_n2_final.json > /tmp/t.md"""
import json
import sys


def main():
    d = json.load(open(sys.argv[1]))
    c, r = d["config"], d["roofline"]
    out = []
    out.append("Headline (`%s`, %d GPU): **%.0f GB/s** of input device resident (compress %.3f ms + decompress %.3f ms, ratio %.2f, "
               "round trip at %.2f of the measured HBM peak %.0f GB/s); dominant kernel `%s` %.3f ms = %.0f GB/s algorithmic = **%.2f** of peak "
               "(compress call %.2f, decompress %.2f)." % (c["workload"], d["n_gpus"], d["value"], c["compress_ms"], c["decompress_ms"],
                                                           c["compression_ratio"], c["roundtrip_frac_of_hbm"], r["peak"], r["kernel"], r["kernel_ms"],
                                                           r["achieved"], r["frac"], r["compress"]["frac"], r["decompress"]["frac"]))
    e = d.get("e2e")
    if e:
        out.append("")
        out.append("End to end through `dcb_compress_host` + `dcb_decompress_host` (host buffers, copies inside): **%.1f GB/s** with %d host "
                   "threads and pinned buffers = %.2f of the box's measured ceiling for this traffic (%.1f GB/s: %s); ONE caller, serial calls: "
                   "%.1f GB/s pinned, %.1f GB/s with pageable (malloc) buffers." % (
                       e["value"], e["host_threads"], e.get("frac_of_ceiling", float("nan")), e.get("ceiling", {}).get("input_gbs", float("nan")),
                       e.get("ceiling", {}).get("how", ""), e.get("serial_pinned_gbs", float("nan")), e.get("serial_pageable_gbs", float("nan"))))
    cb = d.get("cpu_baseline")
    if cb:
        out.append("Reference C code on the host (%s, %d cores): %.2f GB/s." % (cb["kind"], cb["cores"], cb["value"]))
    out.append("")
    out.append("| CT | field | log2 n | absErrorBound | chunk | ratio | compress ms | decompress ms | input GB/s | round trip / HBM | compress / HBM | decompress / HBM | oracle check |")
    out.append("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
    for s in d.get("sweep", []):
        out.append("| %d | %s | %s | %g | %s | %.2f | %.3f | %.3f | %.0f | %.3f | %.3f | %.3f | %s |" % (
            s["ct"], s["field"] + (" (float64)" if s.get("dtype") == "f64" else ""), ("%g" % s["log2n"]), s["eb"], s["chunk"] if s["chunk"] else "one call", s["ratio"], s["compress_ms"],
            s["decompress_ms"], s["input_gbs"], s["roundtrip_frac_of_hbm"], s["compress_frac_of_hbm"], s["decompress_frac_of_hbm"],
            "ok" if s["oracle_check"] else "**FAIL**"))
    if d.get("n4"):
        out.append("")
        out.append("Row N4 (`n4` of the same line; 2^%d floats, chunk 4096, absErrorBound 0.01; Hamming over the CT=5 stream of the noisy field):" % round(d["sweep"][0]["log2n"]))
        out.append("")
        out.append("| kernel | field / size | ratio | compress (encode) ms | decompress (decode) ms | GB/s | / HBM | oracle check |")
        out.append("|---|---|---|---|---|---|---|---|")
        for q in d["n4"]:
            if q["what"] == "bytewise_ct1":
                out.append("| CT=1 byte-wise (`myCompress` / `myDecompress`) | %s | %.2f | %.3f | %.3f | %.0f of input, round trip | %.2f / %.2f | %s |" % (
                    q["field"], q["ratio"], q["compress_ms"], q["decompress_ms"], q["input_gbs"], q["compress_frac_of_hbm"],
                    q["decompress_frac_of_hbm"], "ok" if q["oracle_check"] else "**FAIL**"))
            else:
                out.append("| CT=10 Hamming SECDED, blocks of %d bytes | %d-byte stream, %d blocks | | %.3f | %.3f | %.0f / %.0f | %.2f (encode) | %s |" % (
                    q["block_bytes"], q["stream_bytes"], q["blocks"], q["encode_ms"], q["decode_ms"], q["encode_gbs"], q["decode_gbs"],
                    q["encode_frac_of_hbm"], "ok" if q["oracle_check"] else "**FAIL**"))
    for path in sys.argv[2:]:
        m = json.load(open(path))
        out.append("")
        out.append("%d GPUs (`%s`): %.0f GB/s device resident, end to end %.1f GB/s (%.2f of the ceiling %.1f)." % (
            m["n_gpus"], path.split("/")[-1], m["value"], m.get("e2e", {}).get("value", float("nan")),
            m.get("e2e", {}).get("frac_of_ceiling", float("nan")), m.get("e2e", {}).get("ceiling", {}).get("input_gbs", float("nan"))))
        if m.get("pingpong"):
            out.append("")
            out.append("| pingpong over NCCL (%d GPUs) | raw CT=0 round trip us | CT=5 round trip us | CT=8 round trip us | frame | stream bytes | compress us | decompress us | checks |" % m["n_gpus"])
            out.append("|---|---|---|---|---|---|---|---|---|")
            for p in m["pingpong"]:
                out.append("| %s | %.1f | %.1f | %.1f | %s | %d | %.1f | %.1f | %s |" % (
                    p["case"], p["ct0_round_trip_us"], p["ct5_round_trip_us"], p["ct8_round_trip_us"], p["ct5_frame"], p["ct5_bytes"],
                    p["ct5_compress_us"], p.get("ct5_decompress_us", float("nan")), "ok" if p["ct5_check"] and p["ct8_check"] else "**FAIL**"))
        h = m.get("himeno")
        if h:
            out.append("")
            out.append("Himeno %s, %d sweeps, absErrorBound %g, ms per iteration (sweep + k-face exchange + all-reduce): uncompressed %.3f; "
                       "CT=5 one reference call per face %.3f; CT=5 faces in chunks of 4096 %.3f; CT=5 with the reference's two message rounds "
                       "(round 1 code) %.3f.  Max deviation from the uncompressed run %.3g (bound nn x eb = %.3g)." % (
                           h["grid"], h["nn"], h["abs_error_bound"], h["ct0_ms_per_iter"], h["ct5_one_call_per_face_ms_per_iter"],
                           h["ct5_chunked_4096_ms_per_iter"], h["ct5_two_rounds_r1_ms_per_iter"], h["ct5_one_call_per_face_max_dev_from_ct0"],
                           h["nn"] * h["abs_error_bound"]))
    print("\n".join(out))


if __name__ == "__main__":
    main()
