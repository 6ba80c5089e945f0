python scripts/make_gguf.py --model 8b --quant q4_k_m --layers 2 -o /tmp/8b.gguf > /dev/null
export GGML_BACKEND_PATH=$PWD/super-llama.cpp_b200/libggml-b200.so
GGML_B200_DEBUG=1 oracle/_ref/avx2/llama-bench -m /tmp/8b.gguf -p 0 -n 8 -r 1 -t 16 -fa 1 -ngl 99 2>&1 | grep -E "declined|tg8" | head
GGML_B200_DEBUG=1 oracle/_ref/avx2/llama-bench -m /tmp/8b.gguf -p 512 -n 0 -r 1 -t 16 -fa 1 -ngl 99 2>&1 | grep -E "declined|pp512" | head
