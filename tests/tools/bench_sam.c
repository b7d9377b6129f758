/* bench_sam.c -- TEST TOOL: times the product's host SAM formatting (femhost_sam_read) on mappings
 * produced by the CPU oracle (no GPU needed).  Usage: bench_sam ref.fa index reads.fq e [n_reads] */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>
#include "femhost.h"
#include "../../oracle/fem_oracle.h"
static double now(void){struct timeval tv;gettimeofday(&tv,0);return tv.tv_sec+1e-6*tv.tv_usec;}
int main(int argc,char**argv){
  if(argc<5) return 2;
  int e=atoi(argv[4]); uint32_t n_max=argc>5?(uint32_t)atoi(argv[5]):50000;
  fo_ctx*o=fo_open_reference(argv[1]); fo_load_index(o,argv[2]); fo_set_params(o,0,0,e,1);
  femhost_records ref,reads; femhost_records_init(&ref,0); femhost_records_init(&reads,0);
  femhost_load_all(argv[1],&ref);
  femhost_reader*rd=femhost_reader_open(argv[3]); femhost_reader_next(rd,n_max,&reads); femhost_reader_close(rd);
  uint32_t n=reads.n_records; uint64_t*first=calloc(n+1,8); femgpu_mapping*maps=NULL; size_t cap=0,tot=0;
  fo_mapping*fm=NULL; size_t fcap=0; fo_stats st; memset(&st,0,sizeof st);
  double t=now();
  for(uint32_t i=0;i<n;++i){ uint64_t off=reads.base_off[i]; int l=(int)(reads.base_off[i+1]-off);
    size_t k=fo_map_read(o,reads.bases+off,l,&fm,&fcap,&st);
    if(tot+k>cap){cap=(tot+k)*2+1024;maps=realloc(maps,cap*sizeof*maps);}
    for(size_t j=0;j<k;++j){ femgpu_mapping m; memset(&m,0,sizeof m); m.candidate_position=fm[j].candidate_position; m.end_position_offset=fm[j].end_position_offset; m.direction=fm[j].direction; m.edit_distance=fm[j].edit_distance; maps[tot+j]=m; }
    tot+=k; first[i+1]=tot; }
  fprintf(stderr,"oracle mapped %u reads, %zu mappings in %.2fs\n",n,tot,now()-t);
  femgpu_mapping*work=malloc((tot+1)*sizeof*work);
  femhost_str out={0,0,0};
  for(int rep=0;rep<5;++rep){ memcpy(work,maps,tot*sizeof*work); out.l=0; t=now();
    for(uint32_t i=0;i<n;++i){ uint64_t off=reads.base_off[i]; int l=(int)(reads.base_off[i+1]-off);
      femhost_sam_read(&ref,femhost_record_name(&reads,i),reads.bases+off,reads.quals+off,l,e,work+first[i],NULL,first[i+1]-first[i],&out); }
    double dt=now()-t; fprintf(stderr,"sam_read: %.3fs for %u reads = %.0f ns/read, %zu bytes\n",dt,n,dt*1e9/n,out.l); }
  return 0; }
