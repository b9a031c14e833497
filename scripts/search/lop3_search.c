// Exhaustive search: can the vertical part of the Life rule be done in fewer LOP3 gates?
// Inputs of the back end: two 2-bit row-sum encodings per group + centre bit -> see main().
// Signals are truth tables over NV variables (NV <= 7), stored as 128-bit masks.
#include <stdio.h>
#include <stdlib.h>
#include <stdint.h>
#include <string.h>
typedef unsigned __int128 tt;
static int NV; static tt FULL;
static tt var_tt(int v){ tt r=0; for(int i=0;i<(1<<NV);++i) if((i>>v)&1) r|=(tt)1<<i; return r; }
static inline tt lut3(int lut, tt a, tt b, tt c){ // LOP3 semantics: bit index = a*4+b*2+c
    tt r=0;
    for(int k=0;k<8;++k) if((lut>>k)&1){ tt m=FULL; m&=(k&4)?a:~a; m&=(k&2)?b:~b; m&=(k&1)?c:~c; r|=m; }
    return r&FULL;
}
// is f a function of (a,b,c) on care set? returns lut or -1
static inline int func_of(tt f, tt care, tt a, tt b, tt c){
    int lut=0;
    for(int k=0;k<8;++k){ tt m=care; m&=(k&4)?a:~a; m&=(k&2)?b:~b; m&=(k&1)?c:~c;
        tt fm=f&m; if(fm==0) continue; if(fm==m){ lut|=1<<k; continue;} return -1; }
    return lut;
}
int main(int argc,char**argv){
    // 7 raw variables: a0 b0 c0 (low bits of three row sums), a1 b1 c1 (high bits), x (centre)
    NV=7; FULL=~(tt)0; // 128 bits exactly
    tt a0=var_tt(0),b0=var_tt(1),c0=var_tt(2),a1=var_tt(3),b1=var_tt(4),c1=var_tt(5),x=var_tt(6);
    // target
    tt f=0;
    for(int i=0;i<128;++i){ int s=((i>>0)&1)+((i>>1)&1)+((i>>2)&1)+2*(((i>>3)&1)+((i>>4)&1)+((i>>5)&1)); int xx=(i>>6)&1;
        // row sums include the centre: sum9 = s; next = s==3 || (s==4 && x)
        if(s==3||(s==4&&xx)) f|=(tt)1<<i; }
    // care set: row sum b (middle row) includes x, so (b1,b0) >= x  i.e. if x==1 then b!=0
    tt care=0;
    for(int i=0;i<128;++i){ int b=((i>>1)&1)+2*((i>>4)&1); int xx=(i>>6)&1; if(xx&&b==0) continue; care|=(tt)1<<i; }
    int use_care = argc>1 && atoi(argv[1]);
    if(!use_care) care=FULL;
    int ngates = argc>2? atoi(argv[2]) : 3;
    // front: symmetric encodings of the two groups (fixed 4 gates), choose encoding pair
    // balanced partitions of count {0..3}: P0 parity{1,3}, P1 {2,3}, P2 {0,3}
    int part[3][4]={{0,1,0,1},{0,0,1,1},{1,0,0,1}};
    tt lowp[3],highp[3];
    for(int p=0;p<3;++p){ lowp[p]=0; highp[p]=0;
        for(int i=0;i<128;++i){ int cl=((i>>0)&1)+((i>>1)&1)+((i>>2)&1); int ch=((i>>3)&1)+((i>>4)&1)+((i>>5)&1);
            if(part[p][cl]) lowp[p]|=(tt)1<<i; if(part[p][ch]) highp[p]|=(tt)1<<i; } }
    int pairs[3][2]={{0,1},{0,2},{1,2}};
    for(int el=0;el<3;++el) for(int eh=0;eh<3;++eh){
        tt sig[16]; int ns=0;
        sig[ns++]=lowp[pairs[el][0]]; sig[ns++]=lowp[pairs[el][1]];
        sig[ns++]=highp[pairs[eh][0]]; sig[ns++]=highp[pairs[eh][1]];
        sig[ns++]=x;
        int raw = argc>3? atoi(argv[3]):0;
        if(raw){ sig[ns++]=a0;sig[ns++]=b0;sig[ns++]=c0;sig[ns++]=a1;sig[ns++]=b1;sig[ns++]=c1; }
        int base=ns; long found=0;
        if(ngates==2){
            for(int i=0;i<base;++i)for(int j=i+1;j<base;++j)for(int k=j+1;k<base;++k)for(int l1=0;l1<128;++l1){
                tt g1=lut3(l1*2,sig[i],sig[j],sig[k]);
                for(int p=0;p<base;++p)for(int q=p+1;q<base;++q){ int r=func_of(f,care,g1,sig[p],sig[q]); if(r>=0){ if(!found) printf("enc %d %d: 2 gates: g1=L%02x(%d,%d,%d) out=L%02x(g1,%d,%d)\n",el,eh,l1*2,i,j,k,r,p,q); ++found; } }
            }
        } else {
            for(int i=0;i<base;++i)for(int j=i+1;j<base;++j)for(int k=j+1;k<base;++k)for(int l1=0;l1<128;++l1){
                tt g1=lut3(l1*2,sig[i],sig[j],sig[k]); sig[base]=g1;
                int n2=base+1;
                for(int i2=0;i2<n2;++i2)for(int j2=i2+1;j2<n2;++j2)for(int k2=j2+1;k2<n2;++k2)for(int l2=0;l2<128;++l2){
                    tt g2=lut3(l2*2,sig[i2],sig[j2],sig[k2]); sig[base+1]=g2;
                    int n3=base+2;
                    // gate 3 must use g2
                    for(int p=0;p<n3-1;++p)for(int q=p+1;q<n3-1;++q){
                        if(k2!=base && p!=base && q!=base) continue; // g1 unused
                        int r=func_of(f,care,g2,sig[p],sig[q]);
                        if(r>=0){ if(found<50) printf("enc %d %d: 3 gates: g1=L%02x(%d,%d,%d) g2=L%02x(%d,%d,%d) out=L%02x(g2,%d,%d)\n",el,eh,l1*2,i,j,k,l2*2,i2,j2,k2,r,p,q); ++found; }
                    }
                }
            }
        }
        printf("encoding low=%d high=%d: %ld solutions with %d back gates\n",el,eh,found,ngates); fflush(stdout);
    }
    return 0;
}
