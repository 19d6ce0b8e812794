#!/bin/bash
# Single-GPU evidence batch: zoo sweep, steady-state launch lists, ncu --set full captures of the hot kernels (taken only
# after the same program has exited 0 without ncu), the communication kernels in loopback.  Output: gpurun_out/prof/.
O=gpurun_out/prof; mkdir -p $O
ZOO="LeNet VGG16 ResNet18 ResNet50 PreActResNet18 GoogLeNet DenseNet121 ResNeXt29_2x64d MobileNet MobileNetV2 DPN26 ShuffleNetG2 ShuffleNetV2_1 SENet18 EfficientNetB0 RegNetX_200MF SimpleDLA DLA PNASNetA"
echo "== zoo sweep"; timeout 900 python tools/bench_zoo.py $ZOO > $O/zoo.log 2>&1; grep '"model"' $O/zoo.log | cut -c1-90
echo "== kagg loopback"; timeout 300 python tools/prof_kagg.py 13 45 359 > $O/kagg.log 2>&1 && cp gpurun_out/prof_kagg.json $O/; cat $O/kagg.log | cut -c1-300
for M in ResNet18 MobileNet; do
  echo "== launch list $M"
  ZOO_STEPS=4 timeout 600 python tools/prof_zoo.py $M > $O/eager_$M.log 2>&1 && \
  ZOO_STEPS=4 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_$M.csv python tools/prof_zoo.py $M > $O/ncu_l_$M.log 2>&1
  python tools/launch_summary.py --last-step 4 $O/launches_$M.csv | head -32
done
echo "== ncu full: gemm / pack"; ZOO_STEPS=2 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gemm_tc_kernel|pack_multi_kernel' --launch-skip 60 --launch-count 36 -o $O/gemm_conv python tools/prof_zoo.py ResNet18 > $O/ncu_gemm.log 2>&1; ls -la $O/gemm_conv.ncu-rep
echo "== ncu full: comm kernels"; timeout 600 ncu --set full --clock-control none --import-source on -k regex:'fedavg_kernel|push_copy_kernel|codec_encode_kernel' --launch-skip 3 --launch-count 6 -o $O/kagg python tools/prof_kagg.py 45 > $O/ncu_kagg.log 2>&1; ls -la $O/kagg.ncu-rep
echo "== bench N=1"; timeout 300 python bench.py --steps 20 --warmup 3 > $O/bench_N1.json 2> $O/bench_N1.err; cut -c1-400 $O/bench_N1.json
