vlib=$PWD/chaos-ray-tracing-course-2025_b200/lib/variants/libcrt_b200_libsincos.so
for rep in 1 2; do
echo "lean   c3-gi: $(timeout 300 python tools/render_frames.py c3-gi 3 auto kd 0 2>&1 | tail -1 | cut -c1-60)"
echo "libsin c3-gi: $(CRT_B200_LIB=$vlib timeout 300 python tools/render_frames.py c3-gi 3 auto kd 0 2>&1 | tail -1 | cut -c1-60)"
done
echo "== sah sweep c3-gi"; bash tools/sah_sweep.sh c3-gi
echo "== sah sweep c1"; bash tools/sah_sweep.sh c1
echo "== sah sweep c5"; bash tools/sah_sweep.sh c5
