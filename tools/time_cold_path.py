import time, sys
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import kgx_loader; kgx_loader.load()
from kgx_b200 import kgx
from kgx_b200.abi import MODE_X, MODE_R_INTENDED
from kgx_b200.fleet import Fleet
for cfg,mode in ((4,MODE_X),(3,MODE_R_INTENDED)):
    f=Fleet.config(cfg); t=f.table(mode); s=f.scrape(0)
    g=kgx.Kgx(mode)
    t0=time.time(); g.set_table(t); t1=time.time(); g.ingest(s); t2=time.time(); g.ingest(s); t3=time.time()
    print(cfg, 'set_table %.2fs first ingest (index build) %.3fs steady ingest %.3fs'%(t1-t0,t2-t1,t3-t2), g.stats())
    if mode==MODE_R_INTENDED:
        t0=time.time(); g.set_table(t); t1=time.time(); print('set_table with carry-over %.2fs'%(t1-t0)); g.ingest(s); print(len(g.render()))
