python - <<'PY'
import gzip, os, subprocess, sys, time
sys.path.insert(0, '.')
from oracle import synth
os.makedirs('/tmp/clit', exist_ok=True)
open('/tmp/clit/m.gz','wb').write(gzip.compress(synth.make_ggml(hidden=1024, seed=1234), 1))
synth.write_wav_f32('/tmp/clit/a.wav', synth.make_audio(180*44100, 7))
for conn in ('32', '8', '16'):
  for i in range(2):
    t=time.perf_counter()
    r=subprocess.run(['umx.cpp_b200/umx.cpp.main','/tmp/clit/m.gz','/tmp/clit/a.wav','/tmp/clit/out'],capture_output=True,text=True,env=dict(os.environ,UMX_CLI_TIMES='1',CUDA_DEVICE_MAX_CONNECTIONS=conn))
    print(conn, round(time.perf_counter()-t,2), [l for l in r.stdout.splitlines() if 'phases' in l])
PY
