"""One retrieval scan against a prebuilt CorpusIndex (for ncu captures of the selection GEMM)."""
import sys, torch
sys.path.insert(0, ".")
import two_towers_b200 as tt
Q = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
N = int(sys.argv[2]) if len(sys.argv) > 2 else 1_100_000
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(1)
q = torch.nn.functional.normalize(torch.randn(Q, 128, device=dev, generator=g), dim=1)
docs = torch.nn.functional.normalize(torch.randn(N, 128, device=dev, generator=g), dim=1)
index = tt.CorpusIndex(docs, cosine=True)
v, i = tt.search_topk(q, index, 10)
torch.cuda.synchronize()
print(i[0].tolist())
