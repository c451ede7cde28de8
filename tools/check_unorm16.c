#include <stdio.h>
#include <math.h>
int main(){ const float d=65535.0f; const float r=1.0f/d; int bad=0;
 for(int i=0;i<65536;i++){ float x=(float)i; float q=x*r; float rem=fmaf(-q,d,x); float q2=fmaf(rem,r,q); float ref=x/d; if(q2!=ref){bad++; if(bad<5)printf("%d %a %a\n",i,q2,ref);} }
 printf("bad=%d\n",bad); return 0; }
