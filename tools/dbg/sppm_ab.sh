for v in main r33 r16 r1 main; do
  if [ $v = main ]; then unset B200RT_LIB; else export B200RT_LIB=$PWD/tools/dbg/libb200rt_$v.so; fi
  python tools/bench_sppm.py --no-cpu --iterations 1 --spp 1 --repeat 4 | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1])['gpu']; print('$v', 'alone: photons %.3f ms' % d['ms_photons'], d['global_photons_stored'])"
  python tools/bench_sppm.py --no-cpu --repeat 2 | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1])['gpu']; print('$v', 'full job: total %.1f photons %.1f eye %.1f' % (d['ms_total'], d['ms_photons'], d['ms_eye']))"
done
