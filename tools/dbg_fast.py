import sys, os
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0]=[ROOT+'/snake-slam_b200', ROOT+'/oracle', ROOT+'/tests']
import numpy as np, snakeorb, oracle_py as O
from snakeorb import synthetic as S
img=S.image(1000,752,480)
ex=snakeorb.ORBExtractor(1000,1.2,8,20,7)
k,d=ex.Detect(img)
ref=O.Extractor(1000,1.2,8,20,7,1); k2,d2=ref.detect(img)
got=ex.candidates(0); raw=ref.level_kps(0,0)
want=np.stack([raw['x'],raw['y'],raw['response']],1).astype(np.int32)
print(len(got),len(want))
print('got resp hist', np.bincount(np.minimum(got[:,2],40))[:40])
print('want resp hist', np.bincount(np.minimum(want[:,2],40))[:40])
ws=set(map(tuple,want)); gs=set(map(tuple,got))
print('want-got',len(ws-gs),'got-want',len(gs-ws))
print(sorted(gs-ws)[:20])
print(sorted(ws-gs)[:20])
print('host pixel v', img[16+432,16+3], 'ring0', img[16+432+3,16+3], 'ring4', img[16+432,16+6], 'ring8', img[16+432-3,16+3], 'ring12', img[16+432, 16+0])
