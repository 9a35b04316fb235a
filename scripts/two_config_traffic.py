"""DRAM traffic of serving two STFT configurations (the extractor's (1024, 256, 128) + the encoder's (2048, 512, 80)):
two whole-batch passes against features_and_logmel(..., second=...) in L2-sized chunks.  Run under
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv ...
with MODE=separate | chunked (one mode per process so the launch lists do not mix); CUDA-event time printed without ncu."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixing_style_transfer_b200 as pkg
dev = torch.device("cuda", 0)
B, T = 72, 441000
g = torch.Generator(device=dev).manual_seed(1)
stems = {n: (torch.rand(B, 2, T, device=dev, generator=g) * 2 - 1) * 0.25 for n in pkg.STEMS}
fe = pkg.MixingFeatureExtractor()
enc = pkg.MelSpectrogramPreprocessor(44100, 2048, 512, 80)
mode = os.environ.get("MODE", "chunked")
chunk = int(os.environ.get("CHUNK", "4"))


def run():
    if mode == "separate":
        return fe.features_and_logmel(stems), enc(stems)
    return fe.features_and_logmel(stems, second=(2048, 512, 80), chunk_clips=chunk)


for _ in range(2):
    run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = int(os.environ.get("REPS", "5"))
e0.record()
for _ in range(n):
    run()
e1.record(); torch.cuda.synchronize()
print(f"mode={mode} chunk={chunk}: {e0.elapsed_time(e1) / n:.3f} ms per 72 clips (both configurations)")
