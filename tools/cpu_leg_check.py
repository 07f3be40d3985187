import sys, shutil, json
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
src=open('/root/repo/bench.py').read().split('if __name__ == "__main__":')[0]
b={'__file__':'/root/repo/bench.py','__name__':'bench'}; exec(compile(src,'bench.py','exec'), b)
cfg=b['CONFIGS'][2]
ck,cc=b['make_cpu_sample'](cfg, 1500000)
for core in (0, 5):
    wd=b['_workdir'](); files,nb=b['stage_text_sample'](ck,cc,1,wd)
    R,D,dt=b['cpu_reference_text'](files,cfg,8,wd,[core]); shutil.rmtree(wd)
    print("core",core,"R",R,"dt",round(dt,2), {k:round(v,2) for k,v in b['cpu_reference_text'].stage_s.items()}, flush=True)
