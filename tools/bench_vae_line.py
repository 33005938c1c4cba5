import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
v=d['vae_decode_cupti']; u=d['unet_step_cupti']
print(sys.argv[1], 'img/s', round(d['value'],3), 'unet', round(d['unet_ms_per_step'],3), 'vae busy', v['busy_ms'], {k:round(x['ms'],3) for k,x in v.items() if isinstance(x,dict)}, 'unet gn', round(u['groupnorm_apply']['ms'],3))
