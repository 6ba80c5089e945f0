python -m pytest tests/test_gpu_ops.py tests/test_gpu_backend_ops.py -x -q -k "swiglu or fusions or tensor_core_gemv" 2>&1 | tail -3
for i in 1 2; do
for d in _ab/old .; do (cd $d; python bench.py --no-pp --no-cpu $( [ $d = _ab/old ] || echo --no-llama-bench ) 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$d', d['value'], d['roofline']['frac'], d['roofline']['ms_per_step'])"); done; done
python scripts/make_gguf.py --model 8b --quant q4_k_m -o /tmp/8b.gguf > /dev/null
GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so GGML_B200_GRAPH_STATS=1 oracle/_ref/avx2/llama-bench -m /tmp/8b.gguf -p 0 -n 128 -r 3 -t 16 -fa 1 -ngl 99 2>&1 | grep -E "replays timed|tg128|fusions"
