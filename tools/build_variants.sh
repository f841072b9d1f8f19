# Same-box A/B of kernel variants: builds build/libuz_<name>.so from the CURRENT tree with extra -D flags on one source
# file; select at run time with UZ_LIB_PATH=build/libuz_<name>.so.   usage: tools/build_variants.sh name file.cu -DFLAG=1 ...
set -e
name=$1; file=$2; shift 2
cd "$(dirname "$0")/.."
mkdir -p build/var_$name
for f in unet_zoo_b200/csrc/*.cu; do
  o=build/var_$name/$(basename ${f%.cu}).o
  if [ "$(basename $f)" = "$file" ]; then
    nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC "$@" -c $f -o $o 2>/dev/null &
  else
    cp build/$(basename ${f%.cu}).o $o
  fi
done
wait
nvcc -shared -o build/libuz_$name.so build/var_$name/*.o -cudart static 2>/dev/null
ls -la build/libuz_$name.so
