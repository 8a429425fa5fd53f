export SPLITQ_LIB=$PWD/llm-inference-in-distributed-edge-networks_b200/lib/variants/libsplitq_selt.so
python - <<'PY'
import sys; sys.path.insert(0,'llm-inference-in-distributed-edge-networks_b200')
import torch, splitq
imp=torch.rand(8,4096,device='cuda')
ladder=splitq.ladder_from_fractions([(0.25,2),(0.25,4),(0.25,8)],4096)
for _ in range(3):
    splitq.select_bits(imp,ladder,896); torch.cuda.synchronize()
imp=torch.softmax(torch.randn(8,4096,device='cuda')*2,dim=1)
splitq.select_bits(imp,ladder,896); torch.cuda.synchronize()
PY
