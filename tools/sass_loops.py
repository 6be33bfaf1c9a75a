import re,sys,subprocess
cubin=sys.argv[1]; func=sys.argv[2]
dis=subprocess.run(["nvdisasm","-c",cubin],stdout=subprocess.PIPE,text=True).stdout.split("\n")
inside=False; items=[]
for ln in dis:
    if ln.startswith(".text."):
        inside=func in ln; continue
    if not inside: continue
    m=re.match(r"\s*(\.L_x_\d+):",ln)
    if m: items.append(("L",m.group(1))); continue
    m=re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);",ln)
    if m: items.append(("I",m.group(2)))
labels={}; n=0; ins=[]
for k,v in items:
    if k=="L": labels[v]=n
    else: ins.append(v); n+=1
print("instructions",n)
loops=[]
for i,t in enumerate(ins):
    m=re.search(r"BRA.*?(\.L_x_\d+)",t)
    if m and m.group(1) in labels and labels[m.group(1)]<=i:
        loops.append((labels[m.group(1)],i))
for a,b in sorted(loops,key=lambda x:x[0]-x[1])[:12]:
    body=ins[a:b+1]
    def c(pat): return sum(1 for t in body if re.search(pat,t))
    print("loop %6d..%6d size %5d  BRA %3d BSSY %2d LDG %3d STG %3d LDL/STL %2d SHFL %2d VIMNMX %3d SHF %3d IMAD %3d IADD3/VIADD %3d MOV %3d SEL %3d ISETP %3d CCTL/PREF %2d"%(a,b,b-a+1,c(r"\bBRA"),c("BSSY"),c(r"LDG"),c(r"STG"),c(r"LDL|STL"),c("SHFL"),c("VIMNMX|VIADDMNMX"),c(r"SHF\."),c(r"IMAD"),c(r"IADD3|VIADD "),c(r"\bMOV|IMAD\.MOV"),c(r"\bSEL"),c("ISETP"),c("CCTL|PREFETCH")))
