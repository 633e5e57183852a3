tools/gpu_call.sh r2zc tests smoke bench ncu:bunny:16 ncu:cornell_box:32
