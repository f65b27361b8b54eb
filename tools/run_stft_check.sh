# GPU check of the tensor-core STFT: parity tests, then the configs[1] bench line (kernel ms, roofline fraction)
timeout 300 python -m pytest tests/test_gpu_stft.py -x -q -k "dft256 or stft_matches" 2>&1 | tail -2
b() { SSB_STFT_DFT256=1 timeout 300 python bench.py --workload stft --steps 5 --warmup 3 --no-extras 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('MS', d['ms_per_step'], d['roofline']['frac'], d['clocks'])"; }
b
for v in "$@"; do
  touch speech-separation_b200/csrc/stft_mma.cu; make -C speech-separation_b200/csrc EXTRA="$v" 2>&1 | grep -i " error"; echo "variant $v"; b
done
