python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29530 tools/pcie_bw.py 2>/dev/null | tee gpurun_out/r01s_pcie_bw_n8.json
python tools/pcie_bw.py 2>/dev/null | tee gpurun_out/r01s_pcie_bw_n1.json
CUDA_VISIBLE_DEVICES=5 python tools/pcie_bw.py 2>/dev/null | tee gpurun_out/r01s_pcie_bw_n1_gpu5.json
