# compile one kernel source with ptxas -v:  bash scripts/cc.sh attention_tc.cu
cd /root/repo/ssl4polyp_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -I../../include -I. -Xptxas -v -c $1 -o /tmp/${1%.cu}.o 2>&1 | grep -E "error|warning|registers|spill|Compiling|Function properties" | grep -v "^$" | head -${2:-40}
