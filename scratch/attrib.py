import csv,re,collections,sys
# usage: attrib.py <ncu source csv> <mangled-name-substring> [top]
srccsv, key = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
import os
txt=open(os.environ.get('SASS','/tmp/bytes_sass.txt')).read().split('\n')
start=[i for i,l in enumerate(txt) if l.startswith('.text.') and key in l][0]
line=None; addr2line={}
for l in txt[start+1:]:
    if l.startswith('//---------------------'): break
    m=re.search(r'//## File "([^"]+)", line (\d+)',l)
    if m: line=int(m.group(2)); continue
    m=re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);',l)
    if m: addr2line[int(m.group(1),16)]=line
rows=list(csv.reader(open(srccsv)))
hdr=rows[1]; ix={h:i for i,h in enumerate(hdr)}
body=[r for r in rows[2:] if len(r)==len(hdr) and r[0]!='Address']
base=int(body[0][0],16)
perline=collections.Counter(); samp=collections.Counter(); thr=collections.Counter()
stalls=[h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
st=collections.Counter()
for r in body:
    a=int(r[0],16); ln=addr2line.get(a-base)
    perline[ln]+=int(r[ix['Instructions Executed']] or 0)
    thr[ln]+=int(r[ix['Thread Instructions Executed']] or 0)
    samp[ln]+=int(r[ix['# Samples']] or 0)
    for s_ in stalls:
        try: st[s_]+=int(r[ix[s_]])
        except: pass
tot=sum(perline.values()); ts=sum(samp.values())
src=open(os.environ.get('SRC','/root/repo/triton_client_b200/csrc/bytes.cu')).read().split('\n')
print('total warp inst',tot,'thread inst/warp inst %.1f'%(sum(thr.values())/max(tot,1)),'samples',ts)
S=sum(st.values())
print(' '.join('%s %.0f%%'%(k.replace('stall_',''),100*v/S) for k,v in st.most_common(8)))
for ln,c in sorted(samp.items(), key=lambda kv:-kv[1])[:top]:
    print('%5s %6.2f%% inst %6.2f%% samp lanes %4.1f  %s'%(ln,100*perline[ln]/tot,100*c/ts, thr[ln]/max(perline[ln],1), src[ln-1].strip()[:95] if ln else ''))
