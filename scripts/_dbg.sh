python -m pytest tests/test_gpu_llama_logits.py tests/test_gpu_backend_ops.py -x -q -k "logits or fusions or ROPE" 2>&1 | tail -4
python scripts/make_gguf.py --model 8b --quant q4_k_m -o /tmp/8b.gguf > /dev/null
GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so GGML_B200_GRAPH_STATS=1 oracle/_ref/avx2/llama-bench -m /tmp/8b.gguf -p 512 -n 128 -r 3 -t 16 -fa 1 -ngl 99 2>&1 | grep -E "replays timed|tg128|pp512"
