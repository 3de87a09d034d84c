"""numpy prototype of the local rule behind ub_draw_contours (csrc/annotate.cu), checked against cv2 on random masks:
exterior background by a 4-connected flood from the frame; for every foreground pixel p and every exterior orthogonal
neighbour p + D[d]: step to p + D[d+1] if that is foreground, else to p + D[d+2] if that is; the painted image is the
union of the unit steps' cv2.line(thickness 2) stamps plus the single-point stamp on every border pixel.
Usage: python tools/contour_proto.py      (prints the stamps and the number of mismatching masks: 0)"""
import sys, numpy as np, cv2
from scipy import ndimage
sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
import unet_b200.synth as S

D = [(1,0),(1,-1),(0,-1),(-1,-1),(-1,0),(-1,1),(0,1),(1,1)]   # OpenCV chain codes (dx, dy)

def stamps():
    """pixel sets drawn by cv2 for a unit step in each direction and for a single point (thickness 2)"""
    st = {}
    for d,(dx,dy) in enumerate(D):
        img = np.zeros((11,11), np.uint8)
        cv2.line(img, (5,5), (5+dx,5+dy), 1, 2)
        st[d] = [(x-5,y-5) for y,x in zip(*np.nonzero(img))]
    img = np.zeros((11,11), np.uint8)
    cv2.line(img, (5,5), (5,5), 1, 2)
    st['pt'] = [(x-5,y-5) for y,x in zip(*np.nonzero(img))]
    return st

def exterior(mask):
    H,W = mask.shape
    pad = np.zeros((H+2,W+2), np.uint8); pad[1:-1,1:-1] = mask
    lab,n = ndimage.label(pad==0, structure=[[0,1,0],[1,1,1],[0,1,0]])
    return (lab == lab[0,0])          # padded exterior background

def draw_local(mask, st):
    H,W = mask.shape
    ext = exterior(mask)
    fg = np.zeros((H+2,W+2), bool); fg[1:-1,1:-1] = mask>0
    out = np.zeros((H,W), np.uint8)
    def put(x,y,key):
        for ox,oy in st[key]:
            xx,yy = x+ox,y+oy
            if 0<=xx<W and 0<=yy<H: out[yy,xx]=1
    ys,xs = np.nonzero(mask)
    for y,x in zip(ys,xs):
        px,py = x+1,y+1
        any_crack=False
        for d in (0,2,4,6):
            zx,zy = px+D[d][0], py+D[d][1]
            if not ext[zy,zx]: continue
            any_crack=True
            d1=(d+1)&7; d2=(d+2)&7
            if fg[py+D[d1][1], px+D[d1][0]]: put(x,y,d1)
            elif fg[py+D[d2][1], px+D[d2][0]]: put(x,y,d2)
        if any_crack: put(x,y,'pt')
    return out

def draw_cv(mask):
    cs,_ = cv2.findContours(mask.copy(), cv2.RETR_EXTERNAL, cv2.CHAIN_APPROX_SIMPLE)
    out = np.zeros(mask.shape, np.uint8)
    cv2.drawContours(out, cs, -1, 1, 2)
    return out

if __name__ == "__main__":
    st = stamps()
    for k,v in st.items(): print(k, len(v), sorted(v))
    rng = np.random.default_rng(0)
    bad = 0; n=0
    for trial in range(300):
        h,w = int(rng.integers(5,60)), int(rng.integers(5,80))
        kind = trial % 4
        if kind == 0: m = (rng.random((h,w)) < rng.uniform(0.05,0.9)).astype(np.uint8)
        elif kind == 1: m = S.mask_blobs(h, w, seed=trial, fill=float(rng.uniform(0.05,0.6)))
        elif kind == 2:
            m = np.zeros((h,w),np.uint8)
            for _ in range(int(rng.integers(1,6))):
                cv2.line(m, (int(rng.integers(0,w)),int(rng.integers(0,h))), (int(rng.integers(0,w)),int(rng.integers(0,h))), 1, int(rng.integers(1,4)))
        else:
            m = cv2.dilate((rng.random((h,w)) < 0.03).astype(np.uint8), np.ones((3,3),np.uint8)); m[rng.random((h,w))<0.1]=0
        a = draw_local(m, st); b = draw_cv(m)
        n+=1
        if not np.array_equal(a,b):
            bad+=1
            if bad<=3:
                print("DIFF trial",trial,kind,(h,w), int((a!=b).sum()), "extra", int((a&~b&1).sum()), "missing", int((b&~a&1).sum()))
    print("bad",bad,"of",n)
