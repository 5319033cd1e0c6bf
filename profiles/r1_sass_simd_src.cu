__global__ void k(unsigned* a){ unsigned x=a[0],y=a[1],z=a[2];
 a[3]=__vmaxu4(x,y); a[4]=__vsubus4(x,y); a[5]=__vmaxs2(x,y); a[6]=__vsubss2(x,y); a[7]=__vimax3_s16x2(x,y,z); a[8]=__viaddmax_s16x2(x,y,z);
 a[9]=__vcmpeq4(x,y); a[10]=__vminu4(x,y); a[11]=__vaddus4(x,y); a[12]=__vaddss2(x,y); a[13]=__vimax3_s32(x,y,z); a[14]=__vmaxu2(x,y);a[15]=__vsubus2(x,y);}
