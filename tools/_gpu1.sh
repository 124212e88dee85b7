mkdir -p gpurun_out
for v in "" _u12 _u16 "" _u12 _u16; do echo "== lib$v"; B200FEAT_LIB=$PWD/expressive_speech_translation_b200/lib/libb200feat$v.so timeout 300 python tools/resample_timing.py 1024 2>&1 | grep -v "kaiser\|chain"; done | tee gpurun_out/ab_u.log
