// Search: output gate count of the "row pair" formulation of the Life rule.
//   S = sum of two adjacent row sums (0..6), encoded in 3 bits (q0,q1,q2) by an arbitrary code;
//   A = third row's sum (0..3) in one of the three 2-bit symmetric encodings; m = centre bit.
//   out = [A+S == 3] | m & [A+S == 4];   m = 1 implies S >= 1.
// Question: for which codes does a 3-LOP3 circuit for `out` exist?
// Result: none when the pair contains the centre row (m = 1 implies S >= 1).  With CARE_A=1 in the
// environment the single row is the centre row instead (S = above + below, m = 1 implies A >= 1):
// 36 codes have a 3-gate output stage, but none of those codes can be produced from the two rows
// in 3 gates, so that form does not beat 7 gates either.
// usage: pair_search <part> <nparts>
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
typedef uint64_t tt;
static tt VAR[6];
static inline int func_of(tt f, tt care, tt a, tt b, tt c){
    int lut=0;
    for(int k=0;k<8;++k){ tt m=care; m&=(k&4)?a:~a; m&=(k&2)?b:~b; m&=(k&1)?c:~c;
        tt fm=f&m; if(fm==0) continue; if(fm==m){ lut|=1<<k; continue;} return -1; }
    return lut;
}
static inline tt lut3(int lut, tt a, tt b, tt c){
    tt r=0; for(int k=0;k<8;++k) if((lut>>k)&1){ tt m=~(tt)0; m&=(k&4)?a:~a; m&=(k&2)?b:~b; m&=(k&1)?c:~c; r|=m; } return r; }
// two-gate realisability of F on care over signals sig[0..n)
static int two_gate(tt F, tt care, const tt*sig, int n, int*desc){
    for(int p=0;p<n;++p)for(int q=p+1;q<n;++q){
        tt M[4]; int nc=0; int idx[4];
        for(int k=0;k<4;++k){ tt m=care; m&=(k&2)?sig[p]:~sig[p]; m&=(k&1)?sig[q]:~sig[q]; M[k]=m; tt fm=F&m; if(fm!=0&&fm!=m) idx[nc++]=k; }
        if(nc==0){ desc[0]=-1; desc[1]=p; desc[2]=q; return 1; } // out is a function of p,q alone
        tt dom=0; for(int i=0;i<nc;++i) dom|=M[idx[i]];
        for(int pol=0;pol<(1<<(nc-1));++pol){
            tt g=0; for(int i=0;i<nc;++i){ tt fm=F&M[idx[i]]; if(i>0 && ((pol>>(i-1))&1)) fm=(~F)&M[idx[i]]; g|=fm; }
            for(int x=0;x<n;++x)for(int y=x+1;y<n;++y)for(int z=y+1;z<n;++z){
                int r=func_of(g,dom,sig[x],sig[y],sig[z]);
                if(r>=0){ desc[0]=r; desc[1]=p; desc[2]=q; desc[3]=x; desc[4]=y; desc[5]=z; return 1; }
            }
        }
    }
    return 0;
}
int main(int argc,char**argv){
    int part=argc>1?atoi(argv[1]):0, nparts=argc>2?atoi(argv[2]):1;
    for(int v=0;v<6;++v){ VAR[v]=0; for(int i=0;i<64;++i) if((i>>v)&1) VAR[v]|=(tt)1<<i; }
    // variable order: 0,1 = A code bits; 2,3,4 = q0,q1,q2; 5 = m
    static const int acode[3][4]={{0,1,2,3},{2,1,0,3},{2,0,1,3}}; // value -> 2-bit code (bit0 first signal, bit1 second)
    // class0 (parity,maj): 0->00 1->01(parity) 2->10(maj) 3->11 ; class1 (parity,alleq): 0->(0,1)=2 1->(1,0)=1 2->(0,0)=0 3->3 ; class2 (maj,alleq): 0->(0,1)=2, 1->0, 2->(1,0)=1, 3->3
    long ncase=0, nsol=0;
    int code[7]; code[0]=0;
    for(code[1]=1;code[1]<8;++code[1])
    for(code[2]=1;code[2]<8;++code[2]) if(code[2]!=code[1])
    for(code[3]=1;code[3]<8;++code[3]) if(code[3]!=code[1]&&code[3]!=code[2])
    for(code[4]=1;code[4]<8;++code[4]) if(code[4]!=code[1]&&code[4]!=code[2]&&code[4]!=code[3])
    for(code[5]=1;code[5]<8;++code[5]) if(code[5]!=code[1]&&code[5]!=code[2]&&code[5]!=code[3]&&code[5]!=code[4])
    for(code[6]=code[5];code[6]<8;++code[6]) if(code[6]!=code[1]&&code[6]!=code[2]&&code[6]!=code[3]&&code[6]!=code[4])
    {
        // canonical under bit permutation: require code[1] bits sorted (cheap partial dedupe): skip
        for(int ae=0;ae<3;++ae){
            if((ncase++)%nparts!=part) continue;
            int sOf[8]; for(int c=0;c<8;++c) sOf[c]=-1; for(int s=6;s>=0;--s) sOf[code[s]]=s; // 5/6 may share
            tt F=0, care=0;
            for(int i=0;i<64;++i){ int ac=i&3, qc=(i>>2)&7, m=(i>>5)&1; int A=-1; for(int a=0;a<4;++a) if(acode[ae][a]==ac) A=a;
                int S=sOf[qc]; if(S<0) continue; if(getenv("CARE_A") ? (m&&A==0) : (m&&S==0)) continue; care|=(tt)1<<i;
                // S of a shared code: 5 or 6 both give 0
                int T=A+S; if(T==3||(T==4&&m)) F|=(tt)1<<i; }
            tt sig[8]; for(int v=0;v<6;++v) sig[v]=VAR[v];
            int found=0, desc[6];
            for(int i=0;i<6&&!found;++i)for(int j=i+1;j<6&&!found;++j)for(int k=j+1;k<6&&!found;++k)for(int l=1;l<128&&!found;++l){
                sig[6]=lut3(l*2,sig[i],sig[j],sig[k]);
                if(two_gate(F,care,sig,7,desc)){ found=1;
                    printf("SOL aenc=%d code=%d%d%d%d%d%d%d g1=L%02x(%d,%d,%d) g2=L%02x(%d,%d,%d) out(g2,%d,%d)\n",ae,code[0],code[1],code[2],code[3],code[4],code[5],code[6],l*2,i,j,k,desc[0],desc[3],desc[4],desc[5],desc[1],desc[2]); fflush(stdout); ++nsol; }
            }
        }
    }
    printf("part %d/%d done: %ld cases, %ld with a 3-gate output\n",part,nparts,ncase,nsol);
    return 0;
}
