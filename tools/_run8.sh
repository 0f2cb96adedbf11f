TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 8 --master-port 29521 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/s8_cover.json 2> gpurun_out/s8_cover.err
$TR --nproc-per-node 4 --master-port 29522 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/s4_cover.json 2> gpurun_out/s4_cover.err
$TR --nproc-per-node 8 --master-port 29523 bench.py --gpus 8 --scene cover_100k_motion --width 3840 --height 2160 --spp 512 --steps 1 --warmup 3 > gpurun_out/s8_100k.json 2> gpurun_out/s8_100k.err
for f in s8_cover s4_cover s8_100k; do cut -c1-260 gpurun_out/$f.json; tail -2 gpurun_out/$f.err; done
