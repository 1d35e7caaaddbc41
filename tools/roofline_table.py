"""Per-layer roofline table for CaffeNet b256 from one bench.py JSON line (its `layer_ms_fwd_bwd`, measured with
CUDA events around every layer by sparknet_profile_step).

    python tools/roofline_table.py profiles/r1_bench_caffenet_tf32_v24.json > profiles/r1_roofline_layers_v24.md

Algorithmic work per layer (SURVEY.md App. B shapes, 8(d) accounting):
  contractions  forward 2*MACs; backward 2*MACs for the filter gradient + 2*MACs for the data gradient
                (conv1 has no data gradient: its bottom is the data layer)
  bandwidth     minimal bytes an implementation must move: every input read once, every output written once, fp32
                (max-pool also moves its int32 argmax mask; in-place ReLU that is fused into its producer moves 0)
The time of a layer includes its pack / reduce / bias-gradient helper kernels, so the fraction is for the layer as
built, not for the bare tensor-core kernel (profiles/r1_ncu_gemm_iteration.md has those)."""
import json
import sys

N = 256
# name: (fwd MACs, has data gradient)
GEMM = {
    "conv1": (26986e6, False), "conv2": (57331e6, True), "conv3": (38277e6, True), "conv4": (28708e6, True),
    "conv5": (19139e6, True), "fc6": (9664e6, True), "fc7": (4295e6, True), "fc8": (1049e6, True),
}
E = 4  # bytes per element


def elts(c, h, w=None):
    return N * c * h * (w if w is not None else h)


# name: (fwd bytes, bwd bytes, note)
BW = {
    # max pool: fwd reads x, writes y + int32 mask; bwd reads dy + mask (+ the POOLED output as the gate of the ReLU folded
    # in: the element that won a window is > 0 exactly when the pooled value is), writes dx
    "pool1": (E * (elts(96, 55) + 2 * elts(96, 27)), E * (3 * elts(96, 27) + elts(96, 55)), "bwd includes relu1's backward (gate = pooled output)"),
    "pool2": (E * (elts(256, 27) + 2 * elts(256, 13)), E * (3 * elts(256, 13) + elts(256, 27)), "bwd includes relu2's backward"),
    "pool5": (E * (elts(256, 13) + 2 * elts(256, 6)), E * (3 * elts(256, 6) + elts(256, 13)), "bwd includes relu5's backward"),
    # LRN across channels without a stored scale tensor: fwd reads x writes y; bwd reads x, y, dy writes dx
    # (round 2: the forward kernel also writes the next convolution's packed input, 4 B per element more;
    #  the backward kernel recomputes y and reads x, dy only: 12 B per element)
    "norm1": (E * 3 * elts(96, 27), E * 3 * elts(96, 27), "no scale tensor kept; forward also writes conv2's packed input"),
    "norm2": (E * 3 * elts(256, 13), E * 3 * elts(256, 13), "no scale tensor kept; forward also writes conv3's packed input"),
    # dropout: fwd reads x writes y + byte mask; bwd reads dy + mask writes dx
    "drop6": (N * 4096 * (2 * E + 1), N * 4096 * (2 * E + 1), ""),
    "drop7": (N * 4096 * (2 * E + 1), N * 4096 * (2 * E + 1), ""),
}
PARAMS = 60965224


def main():
    path = sys.argv[1]
    d = json.loads(open(path).read().strip().splitlines()[-1])
    lm = d["layer_ms_fwd_bwd"]
    tf32_peak = d["roofline"]["peak"]
    hbm_peak = d["roofline_hbm"]["peak"]
    print(f"# Per-layer roofline, CaffeNet b256 TF32 on one B200 — from `{path}`\n")
    print(f"Step {d['ms_per_step']:.3f} ms ({d['value']:.0f} images/s).  Tensor peak {tf32_peak:.0f} TFLOP/s (TF32 cuBLAS 8192^3, "
          f"measured in the same run); HBM peak {hbm_peak:.0f} GB/s (MEASURED_PEAKS.json).  Times are CUDA-event times "
          "around each layer of one iteration (`sparknet_profile_step`), helper kernels (packs, slice reductions, bias "
          "gradients) included.\n")
    print("## Contractions (tensor-bound)\n")
    print("| layer | fwd GFLOP | fwd ms | fwd TFLOP/s | frac | bwd GFLOP | bwd ms | bwd TFLOP/s | frac |")
    print("|---|---|---|---|---|---|---|---|---|")
    tf, tb, msf, msb = 0.0, 0.0, 0.0, 0.0
    for name, (macs, dgrad) in GEMM.items():
        f_ms, b_ms = lm[name]
        gf, gb = 2 * macs / 1e9, 2 * macs * (2 if dgrad else 1) / 1e9
        tf += gf; tb += gb; msf += f_ms; msb += b_ms
        print(f"| {name} | {gf:.1f} | {f_ms:.3f} | {gf / f_ms:.0f} | {gf / f_ms / tf32_peak:.2f} | {gb:.1f} | {b_ms:.3f} | "
              f"{gb / b_ms:.0f} | {gb / b_ms / tf32_peak:.2f} |")
    print(f"| **all** | {tf:.1f} | {msf:.3f} | {tf / msf:.0f} | {tf / msf / tf32_peak:.2f} | {tb:.1f} | {msb:.3f} | {tb / msb:.0f} | "
          f"{tb / msb / tf32_peak:.2f} |")
    print("\nfc6-fc8 are weight-streaming at batch 256 (234 MB of weights read per direction, written once as gradient): "
          "their bound is HBM / L2, not the tensor pipe.\n")
    print("## Bandwidth-bound layers (HBM)\n")
    print("| layer | fwd MB | fwd ms | fwd GB/s | frac | bwd MB | bwd ms | bwd GB/s | frac | note |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    for name, (fb, bb, note) in BW.items():
        f_ms, b_ms = lm[name]
        fcell = f"{fb / 1e6:.1f} | {f_ms:.3f} | {fb / 1e6 / f_ms:.0f} | {fb / 1e6 / f_ms / hbm_peak:.2f}" if fb else "0 | - | - | -"
        print(f"| {name} | {fcell} | {bb / 1e6:.1f} | {b_ms:.3f} | {bb / 1e6 / b_ms:.0f} | {bb / 1e6 / b_ms / hbm_peak:.2f} | {note} |")
    sgd_ms = d["roofline_hbm"]["sgd_update_ms"]
    sgd_b = 24.0 * PARAMS
    print(f"| sgd_update | - | - | - | - | {sgd_b / 1e6:.1f} | {sgd_ms:.3f} | {sgd_b / 1e6 / sgd_ms:.0f} | {sgd_b / 1e6 / sgd_ms / hbm_peak:.2f} | "
          "24 B per parameter algorithmic; the kernel also writes the 4 B tf32 shadow |")
    print("\nSmall layers (dropout on 1 M elements, softmax on 256 x 1000) are launch-latency bound, not bandwidth bound.")


if __name__ == "__main__":
    main()
