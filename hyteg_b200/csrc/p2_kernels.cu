// P2 elementwise operators, fused and matrix-free (sm_100a): per micro-tet the blended element matrix is never formed -- the kernel
// evaluates y_e = A_e u_e quadrature point by quadrature point from the geometry and scatters the ten results.
//
// Replaces P2ElementwiseOperator< Form >::gemv (reference src/hyteg/elementwiseoperators/P2ElementwiseOperator.cpp:109-320: per micro-cell
// assembleLocalElementMatrix3D -> form.integrateAll (forms/form_hyteg_generated/p2/p2_div_k_grad_blending_q4.cpp:848-,
// p2_epsilonvar_blending_q2.cpp) + localMatrixVectorMultiply3D (:67-107) + communicateAdditively (:229-233)) and the hyteg_operators
// kernels BASELINE config 3 / 4 name (P2ElementwiseDivKGrad...IcosahedralShellMap, P2ElementwiseEpsilon..._i_j; their source is absent from
// the reference tree, so the in-tree forms are the definition).
//
// At a quadrature point q with barycentric coordinates lambda (constant tables below, read off the generated forms):
//     M = DF(x_q) J_aff,   s = w_q k_q |det M|,
//     grad u = M^-T sum_j u_j gradhat phi_j,   f = s grad u   (epsilon block (c,d): f = s ( delta_cd grad u + (grad u)_c e_d )),
//     y_i += gradhat phi_i . ( M^-1 f ).
// With the P2 basis in barycentric form (phi_i = lambda_i (2 lambda_i - 1), phi_ab = 4 lambda_a lambda_b) the two sums over the ten basis
// functions collapse to a handful of operations on the four barycentric directions, so the geometry (DF, the 3x3 inverse) is the
// bulk of the ~350 flop per quadrature point: an FP64-pipe-bound kernel.
//
// Scatter: FP64 atomics (RED.ADD.F64) into zeroed accumulators with the function's layout; a second kernel adds the partial sums of
// the copies of every interface DoF (ascending macro-cell id) and applies DoFType selection and beta -- the role of
// communicateAdditively.  Kernels are written one-item-per-thread without warp intrinsics so that the TEST-ONLY emulation build
// (hb_compat.h) can run them on the CPU.
#include "p2_kernels.h"

#include <algorithm>
#include <cmath>

#include "hb_device.h"

#ifdef HB_EMUL
static inline void hb_atomic_add( double* p, double v )
{
   *p += v;
}
#else
__device__ __forceinline__ void hb_atomic_add( double* p, double v )
{
   atomicAdd( p, v );
}
#endif

// ---- constant tables ---------------------------------------------------------------------------------------------------------
// vertices of the six micro-cell types WHITE_UP, BLUE_UP, GREEN_UP, WHITE_DOWN, BLUE_DOWN, GREEN_DOWN as offsets from the cell's logical
// index (reference volumedofspace/CellDoFIndexing.hpp:155-198)
HB_CONSTANT int8_t hb_p2_vert[6][4][3] = { { { 0, 0, 0 }, { 1, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 } }, { { 1, 0, 0 }, { 1, 1, 0 }, { 0, 1, 0 }, { 1, 0, 1 } },
                                           { { 1, 0, 0 }, { 0, 1, 0 }, { 1, 0, 1 }, { 0, 0, 1 } }, { { 1, 1, 0 }, { 1, 1, 1 }, { 0, 1, 1 }, { 1, 0, 1 } },
                                           { { 1, 0, 1 }, { 0, 1, 1 }, { 0, 0, 1 }, { 0, 1, 0 } }, { { 0, 1, 0 }, { 1, 1, 0 }, { 1, 0, 1 }, { 0, 1, 1 } } };
// a micro-cell of type t exists at (x,y,z) iff x + y + z <= N - slack[t]
HB_CONSTANT int8_t hb_p2_slack[6] = { 1, 2, 2, 3, 2, 2 };
// the six edges of a micro-cell in the element's local order (2,3) (1,3) (1,2) (0,3) (0,2) (0,1) (EdgeDoFIndexing.hpp:1365-1396) as
// { orientation, dx, dy, dz }: derived from the vertex table with calcEdgeDoFOrientation / calcEdgeDoFIndex (:67-166)
HB_CONSTANT int8_t hb_p2_edge[6][6][4] = {
    { { 5, 0, 0, 0 }, { 4, 0, 0, 0 }, { 3, 0, 0, 0 }, { 2, 0, 0, 0 }, { 1, 0, 0, 0 }, { 0, 0, 0, 0 } },
    { { 6, 0, 0, 0 }, { 5, 1, 0, 0 }, { 0, 0, 1, 0 }, { 2, 1, 0, 0 }, { 3, 0, 0, 0 }, { 1, 1, 0, 0 } },
    { { 0, 0, 0, 1 }, { 5, 0, 0, 0 }, { 6, 0, 0, 0 }, { 4, 0, 0, 0 }, { 2, 1, 0, 0 }, { 3, 0, 0, 0 } },
    { { 3, 0, 0, 1 }, { 1, 1, 0, 1 }, { 0, 0, 1, 1 }, { 5, 1, 0, 0 }, { 4, 0, 1, 0 }, { 2, 1, 1, 0 } },
    { { 5, 0, 0, 0 }, { 2, 0, 1, 0 }, { 1, 0, 0, 1 }, { 6, 0, 0, 0 }, { 0, 0, 0, 1 }, { 3, 0, 0, 1 } },
    { { 3, 0, 0, 1 }, { 4, 0, 1, 0 }, { 5, 1, 0, 0 }, { 2, 0, 1, 0 }, { 6, 0, 0, 0 }, { 0, 0, 1, 0 } } };
HB_CONSTANT int8_t hb_p2_pair[6][2] = { { 2, 3 }, { 1, 3 }, { 1, 2 }, { 0, 3 }, { 0, 2 }, { 0, 1 } };
// end points of an edge DoF per orientation (EdgeDoFIndexing.hpp:170-209)
HB_CONSTANT int8_t hb_p2_ends[7][2][3] = { { { 0, 0, 0 }, { 1, 0, 0 } }, { { 0, 0, 0 }, { 0, 1, 0 } }, { { 0, 0, 0 }, { 0, 0, 1 } }, { { 1, 0, 0 }, { 0, 1, 0 } },
                                           { { 1, 0, 0 }, { 0, 0, 1 } }, { { 0, 1, 0 }, { 0, 0, 1 } }, { { 0, 1, 0 }, { 1, 0, 1 } } };

// Xiao-Gimbutas degree-4 rule, 11 points (barycentric w.r.t. vertices 0..3; weights sum to 1/6): p2_div_k_grad_blending_q4.cpp
HB_CONSTANT double hb_p2_q4[11][5] = {
    { 0.17469405869723054, 0.040490506727590428, 0.01356070187980288, 0.77125473269537614, 0.006541848487473325 },
    { 0.081404918402859225, 0.75250850700965499, 0.068099370938206658, 0.097987203649279112, 0.0092122281926561474 },
    { 0.30444840243449711, 0.24027666492807259, 0.12680172591539199, 0.32847320672203839, 0.032235135341605742 },
    { 0.74122888209362259, 0.067223294893383398, 0.035183929773598722, 0.15636389323939531, 0.0092322998119293929 },
    { 0.053341239535745266, 0.41926631387951302, 0.047781435559086663, 0.47961101102565512, 0.0099888641910932524 },
    { 0.43295349048135578, 0.45076587609127677, 0.059456616299433829, 0.056824017127933668, 0.011578327656272558 },
    { 0.53800720391618573, 0.12941137378891041, 0.33019041483746447, 0.0023910074574393651, 0.012693785874259723 },
    { 0.008991260093335729, 0.1215419913339278, 0.30649398842969028, 0.56297276014304609, 0.013237780011337548 },
    { 0.10660417256199362, 0.097204644587583267, 0.68439041545304002, 0.11180076739738309, 0.017744672359248346 },
    { 0.32923295974264682, 0.029569495206479609, 0.31790356021339461, 0.32329398483747901, 0.018372372071416277 },
    { 0.10384411641099309, 0.43271023904776862, 0.35382323920929709, 0.1096224053319412, 0.025829352669374347 } };
// 4-point degree-2 rule of p2_epsilonvar_blending_q2.cpp
#define HB_QA 0.13819660112501059
#define HB_QB 0.58541019662496829
HB_CONSTANT double hb_p2_q2[4][5] = { { HB_QA, HB_QA, HB_QA, HB_QB, 0.041666666666666657 },
                                      { HB_QA, HB_QA, HB_QB, HB_QA, 0.041666666666666657 },
                                      { HB_QA, HB_QB, HB_QA, HB_QA, 0.041666666666666657 },
                                      { HB_QB, HB_QA, HB_QA, HB_QA, 0.041666666666666657 } };

int hb_p2_num_quad_points( int form )
{
   return hb_p2_form_q2( form ) ? 4 : 11;
}

// ---- geometry ----------------------------------------------------------------------------------------------------------------

// IcosahedralShellMap::evalF / evalDF in derived form (p2_kernels.h: HbP2CellGeom); DF row-major
__host__ __device__ __forceinline__ void hb_shell_map( const HbP2CellGeom& g, const double x[3], double F[3], double DF[9] )
{
   const double vx  = ( x[0] - g.ray[0] ) * g.n[0] + ( x[1] - g.ray[1] ) * g.n[1] + ( x[2] - g.ray[2] ) * g.n[2];
   const double r   = g.radRay + vx * g.drOverVolT;
   const double n2  = x[0] * x[0] + x[1] * x[1] + x[2] * x[2];
   const double inv = 1.0 / sqrt( n2 );
   const double a   = r * inv;            // r / |x|
   const double b   = r * inv * inv * inv; // r / |x|^3
   F[0]             = a * x[0];
   F[1]             = a * x[1];
   F[2]             = a * x[2];
#pragma unroll
   for ( int i = 0; i < 3; i++ )
#pragma unroll
      for ( int j = 0; j < 3; j++ )
         DF[3 * i + j] = ( i == j ? a : 0.0 ) + x[i] * ( g.n[j] * g.drOverVolT * inv - b * x[j] );
}

// M^-1 (row-major) and det M
__host__ __device__ __forceinline__ double hb_inv3( const double M[9], double I[9] )
{
   const double c00 = M[4] * M[8] - M[5] * M[7], c01 = M[5] * M[6] - M[3] * M[8], c02 = M[3] * M[7] - M[4] * M[6];
   const double det = M[0] * c00 + M[1] * c01 + M[2] * c02;
   const double id  = 1.0 / det;
   I[0]             = c00 * id;
   I[1]             = ( M[2] * M[7] - M[1] * M[8] ) * id;
   I[2]             = ( M[1] * M[5] - M[2] * M[4] ) * id;
   I[3]             = c01 * id;
   I[4]             = ( M[0] * M[8] - M[2] * M[6] ) * id;
   I[5]             = ( M[2] * M[3] - M[0] * M[5] ) * id;
   I[6]             = c02 * id;
   I[7]             = ( M[1] * M[6] - M[0] * M[7] ) * id;
   I[8]             = ( M[0] * M[4] - M[1] * M[3] ) * id;
   return det;
}

// computational position of micro-vertex (x,y,z) of a macro-cell
__host__ __device__ __forceinline__ void hb_p2_vertex_pos( const HbP2CellGeom& g, double invN, int x, int y, int z, double p[3] )
{
#pragma unroll
   for ( int d = 0; d < 3; d++ )
      p[d] = g.v[0][d] + ( x * ( g.v[1][d] - g.v[0][d] ) + y * ( g.v[2][d] - g.v[0][d] ) + z * ( g.v[3][d] - g.v[0][d] ) ) * invN;
}

// ---- the element kernel --------------------------------------------------------------------------------------------------------
// NC = 1: scalar forms (div-k-grad, one epsilon block, div, divT); NC = 3: the whole epsilon operator on a vector function (the
// geometry of a quadrature point is evaluated once for all nine blocks).

template < int NC >
__global__ void hb_p2_elem_kernel( const HbP2View pv, const HbP2ElemArgs a )
{
   const int perCell = ( pv.nAnchors + blockDim.x - 1 ) / blockDim.x;
   const int cell    = blockIdx.x / perCell;
   const int ai      = ( blockIdx.x - cell * perCell ) * blockDim.x + threadIdx.x;
   if ( cell >= pv.nCells || ai >= pv.nAnchors )
      return;
   const int          N = pv.N;
   const HbBndPoint   p = pv.anchors[ai];
   const int          x = p.x, y = p.y, z = p.z;
   const HbP2CellGeom g = pv.geom[cell];
   const size_t       vb = (size_t) cell * pv.vCellSize, eb = (size_t) cell * pv.eCellSize;
   const double       invN = 1.0 / N;
   const bool         q2   = hb_p2_form_q2( a.form );
   const int          nq   = q2 ? 4 : 11;
   const bool         p1In = a.form == HB_P2_DIVT, p1Out = a.form == HB_P2_DIV;
   const bool         stokes = NC == 3 && a.form == HB_P2_STOKES;
   for ( int t = 0; t < 6; t++ )
   {
      if ( x + y + z > N - hb_p2_slack[t] )
         continue;
      // local DoFs
      int    vi[4], ei[6];
      double X[4][3], u[NC][10], kk[10], pk[4], yp[4];
#pragma unroll
      for ( int k = 0; k < 4; k++ )
      {
         const int vx = x + hb_p2_vert[t][k][0], vy = y + hb_p2_vert[t][k][1], vz = z + hb_p2_vert[t][k][2];
         vi[k]        = hb_index( N, vx, vy, vz );
         hb_p2_vertex_pos( g, invN, vx, vy, vz, X[k] );
#pragma unroll
         for ( int c = 0; c < NC; c++ )
            u[c][k] = a.srcV[c][vb + vi[k]];
         pk[k] = stokes ? a.srcP[vb + vi[k]] : 0.0;
         yp[k] = 0.0;
      }
#pragma unroll
      for ( int k = 0; k < 6; k++ )
      {
         ei[k] = hb_edge_index( N, hb_p2_edge[t][k][0], x + hb_p2_edge[t][k][1], y + hb_p2_edge[t][k][2], z + hb_p2_edge[t][k][3] );
#pragma unroll
         for ( int c = 0; c < NC; c++ )
            u[c][4 + k] = p1In ? 0.0 : a.srcE[c][eb + ei[k]];
      }
      if ( a.kV )
      {
#pragma unroll
         for ( int k = 0; k < 4; k++ )
            kk[k] = a.kV[vb + vi[k]];
#pragma unroll
         for ( int k = 0; k < 6; k++ )
            kk[4 + k] = a.kE[eb + ei[k]];
      }
      // affine Jacobian, columns = edge vectors of the micro-tet
      double J[9];
#pragma unroll
      for ( int d = 0; d < 3; d++ )
      {
         J[3 * d + 0] = X[1][d] - X[0][d];
         J[3 * d + 1] = X[2][d] - X[0][d];
         J[3 * d + 2] = X[3][d] - X[0][d];
      }
      double yv[NC][10];
#pragma unroll
      for ( int c = 0; c < NC; c++ )
#pragma unroll
         for ( int k = 0; k < 10; k++ )
            yv[c][k] = 0.0;
      const double* kq = a.kq ? a.kq + ( ( (size_t) cell * 6 + t ) * pv.nAnchors + ai ) * nq : nullptr;
      for ( int q = 0; q < nq; q++ )
      {
         const double* Q = q2 ? hb_p2_q2[q] : hb_p2_q4[q];
         const double  lam[4] = { Q[0], Q[1], Q[2], Q[3] };
         double        M[9], Mi[9];
         if ( pv.blending )
         {
            double xq[3], F[3], DF[9];
#pragma unroll
            for ( int d = 0; d < 3; d++ )
               xq[d] = lam[0] * X[0][d] + lam[1] * X[1][d] + lam[2] * X[2][d] + lam[3] * X[3][d];
            hb_shell_map( g, xq, F, DF );
#pragma unroll
            for ( int i = 0; i < 3; i++ )
#pragma unroll
               for ( int j = 0; j < 3; j++ )
                  M[3 * i + j] = DF[3 * i] * J[j] + DF[3 * i + 1] * J[3 + j] + DF[3 * i + 2] * J[6 + j];
         }
         else
         {
#pragma unroll
            for ( int i = 0; i < 9; i++ )
               M[i] = J[i];
         }
         const double det = hb_inv3( M, Mi );
         // coefficient
         double kval = 1.0;
         if ( kq )
            kval = kq[q];
         else if ( a.kV )
         {
            kval = 0.0;
#pragma unroll
            for ( int k = 0; k < 4; k++ )
               kval += kk[k] * ( lam[k] * ( 2.0 * lam[k] - 1.0 ) );
#pragma unroll
            for ( int k = 0; k < 6; k++ )
               kval += kk[4 + k] * ( 4.0 * lam[hb_p2_pair[k][0]] * lam[hb_p2_pair[k][1]] );
         }
         const double s = Q[4] * kval * fabs( det );
         if ( a.diagonal )
         {
            // A_ii = sum_q s |grad phi_i|^2 (+ s (d_c phi_i)^2 for the epsilon block (c,c) / component c of the vector operator)
#pragma unroll
            for ( int k = 0; k < 10; k++ )
            {
               double gh[3]; // gradhat phi_k
               if ( k < 4 )
               {
                  const double f4 = 4.0 * lam[k] - 1.0;
                  gh[0] = k == 0 ? -f4 : ( k == 1 ? f4 : 0.0 );
                  gh[1] = k == 0 ? -f4 : ( k == 2 ? f4 : 0.0 );
                  gh[2] = k == 0 ? -f4 : ( k == 3 ? f4 : 0.0 );
               }
               else
               {
                  const int pa = hb_p2_pair[k - 4][0], pb = hb_p2_pair[k - 4][1];
#pragma unroll
                  for ( int d = 0; d < 3; d++ )
                  {
                     const double ga = pa == 0 ? -1.0 : ( pa == d + 1 ? 1.0 : 0.0 ), gb = pb == 0 ? -1.0 : ( pb == d + 1 ? 1.0 : 0.0 );
                     gh[d]           = 4.0 * ( lam[pa] * gb + lam[pb] * ga );
                  }
               }
               double gp[3];
#pragma unroll
               for ( int d = 0; d < 3; d++ )
                  gp[d] = Mi[d] * gh[0] + Mi[3 + d] * gh[1] + Mi[6 + d] * gh[2];
               const double n2 = gp[0] * gp[0] + gp[1] * gp[1] + gp[2] * gp[2];
#pragma unroll
               for ( int cc = 0; cc < NC; cc++ )
               {
                  const int comp = NC == 3 ? cc : a.compC;
                  yv[cc][k] += s * ( n2 + ( ( NC == 3 || a.form == HB_P2_EPSILON ) ? gp[comp] * gp[comp] : 0.0 ) );
               }
            }
            continue;
         }
         // physical gradient of every component: grad u = M^-T gradhat u, gradhat u from the coefficients c[k] of grad lambda_k
         double gu[NC][3];
         double pval = 0.0; // HB_P2_DIVT: the P1 source at the quadrature point
#pragma unroll
         for ( int cc = 0; cc < NC; cc++ )
         {
            double c[4];
            if ( p1In )
            {
#pragma unroll
               for ( int k = 0; k < 4; k++ )
                  pval += u[cc][k] * lam[k];
               gu[cc][0] = gu[cc][1] = gu[cc][2] = 0.0;
               continue;
            }
#pragma unroll
            for ( int k = 0; k < 4; k++ )
               c[k] = u[cc][k] * ( 4.0 * lam[k] - 1.0 );
#pragma unroll
            for ( int k = 0; k < 6; k++ )
            {
               const int pa = hb_p2_pair[k][0], pb = hb_p2_pair[k][1];
               c[pa] += 4.0 * u[cc][4 + k] * lam[pb];
               c[pb] += 4.0 * u[cc][4 + k] * lam[pa];
            }
            const double th[3] = { c[1] - c[0], c[2] - c[0], c[3] - c[0] }; // grad lambda_0 = -(1,1,1), grad lambda_k = e_k
#pragma unroll
            for ( int d = 0; d < 3; d++ )
               gu[cc][d] = Mi[d] * th[0] + Mi[3 + d] * th[1] + Mi[6 + d] * th[2];
         }
         double pq = 0.0; // HB_P2_STOKES: - w |det M| p(x_q), added to the diagonal of the stress
         if ( stokes )
         {
            const double s0   = Q[4] * fabs( det );
            const double divu = gu[0][0] + gu[NC > 1 ? 1 : 0][NC > 1 ? 1 : 0] + gu[NC > 2 ? 2 : 0][NC > 2 ? 2 : 0];
#pragma unroll
            for ( int k = 0; k < 4; k++ )
            {
               pq -= s0 * pk[k] * lam[k];
               yp[k] -= s0 * lam[k] * divu;
            }
         }
         if ( p1Out )
         {
            // y_i (P1) -= s lambda_i d_c u
            const double v = -s * gu[0][a.compC];
#pragma unroll
            for ( int k = 0; k < 4; k++ )
               yv[0][k] += lam[k] * v;
            continue;
         }
#pragma unroll
         for ( int cc = 0; cc < NC; cc++ )
         {
            double f[3];
            if ( NC == 3 ) // row cc of mu ( grad u + grad u^T )
            {
#pragma unroll
               for ( int d = 0; d < 3; d++ )
                  f[d] = s * ( gu[cc][d] + gu[d < NC ? d : 0][cc] ) + ( d == cc ? pq : 0.0 );
            }
            else if ( a.form == HB_P2_EPSILON )
            {
#pragma unroll
               for ( int d = 0; d < 3; d++ )
                  f[d] = ( a.compC == a.compD ? s * gu[0][d] : 0.0 ) + ( d == a.compD ? s * gu[0][a.compC] : 0.0 );
            }
            else if ( p1In )
            {
#pragma unroll
               for ( int d = 0; d < 3; d++ )
                  f[d] = d == a.compC ? -s * pval : 0.0;
            }
            else
            {
#pragma unroll
               for ( int d = 0; d < 3; d++ )
                  f[d] = s * gu[0][d];
            }
            // hh = M^-1 f ; y_i += gradhat phi_i . hh
            double hh[3];
#pragma unroll
            for ( int d = 0; d < 3; d++ )
               hh[d] = Mi[3 * d] * f[0] + Mi[3 * d + 1] * f[1] + Mi[3 * d + 2] * f[2];
            const double h[4] = { -( hh[0] + hh[1] + hh[2] ), hh[0], hh[1], hh[2] };
#pragma unroll
            for ( int k = 0; k < 4; k++ )
               yv[cc][k] += ( 4.0 * lam[k] - 1.0 ) * h[k];
#pragma unroll
            for ( int k = 0; k < 6; k++ )
            {
               const int pa = hb_p2_pair[k][0], pb = hb_p2_pair[k][1];
               yv[cc][4 + k] += 4.0 * ( lam[pa] * h[pb] + lam[pb] * h[pa] );
            }
         }
      }
#pragma unroll
      for ( int c = 0; c < NC; c++ )
      {
#pragma unroll
         for ( int k = 0; k < 4; k++ )
            hb_atomic_add( a.accV[c] + vb + vi[k], yv[c][k] );
         if ( !p1Out )
         {
#pragma unroll
            for ( int k = 0; k < 6; k++ )
               hb_atomic_add( a.accE[c] + eb + ei[k], yv[c][4 + k] );
         }
      }
      if ( stokes )
      {
#pragma unroll
         for ( int k = 0; k < 4; k++ )
            hb_atomic_add( a.accP + vb + vi[k], yp[k] );
      }
   }
}

void hb_p2_launch_elements( const HbP2View& pv, const HbP2ElemArgs& a, cudaStream_t st )
{
   const int block   = 128;
   const int perCell = ( pv.nAnchors + block - 1 ) / block;
   if ( pv.nCells * perCell <= 0 )
      return;
   if ( a.form == HB_P2_EPSILON_VECTOR || a.form == HB_P2_STOKES )
      HB_LAUNCH( hb_p2_elem_kernel< 3 >, pv.nCells * perCell, block, st, pv, a );
   else
      HB_LAUNCH( hb_p2_elem_kernel< 1 >, pv.nCells * perCell, block, st, pv, a );
}

// ---- finalize: interior entries, then the interface groups ------------------------------------------------------------------------

__global__ void hb_p2_finalize_interior_kernel( const HbP2View pv, const HbP2FinalizeArgs a )
{
   const long long nV = (long long) pv.nCells * pv.vCellSize, nE = (long long) pv.nCells * pv.eCellSize;
   const long long i  = (long long) blockIdx.x * blockDim.x + threadIdx.x;
   if ( i >= nV + nE )
      return;
   const bool      edge  = i >= nV;
   const long long j     = edge ? i - nV : i;
   const int       csz   = edge ? pv.eCellSize : pv.vCellSize;
   const int       cell  = (int) ( j / csz );
   const int       local = (int) ( j - (long long) cell * csz );
   const int       mask  = edge ? pv.emask[local] : pv.vmask[local];
   if ( mask != 0 )
      return; // interface entries belong to the groups
   if ( !( ( a.selMask >> pv.cells[cell].region[0].bflag ) & 1 ) )
      return;
   double*       d   = edge ? a.dstE : a.dstV;
   const double* acc = edge ? a.accE : a.accV;
   if ( d )
      d[j] = ( a.beta == 0.0 ? 0.0 : a.beta * d[j] ) + acc[j];
}

__global__ void hb_p2_finalize_groups_kernel( const HbP2View pv, const HbP2FinalizeArgs a )
{
   const long long gidx = (long long) blockIdx.x * blockDim.x + threadIdx.x;
   if ( gidx >= a.nGroups )
      return;
   if ( !( ( a.selMask >> a.grpFlag[gidx] ) & 1 ) )
      return;
   const long long EDGE = 1ll << 62;
   const long long e0 = a.grpStart[gidx], e1 = a.grpStart[gidx + 1];
   if ( ( a.grpEntry[e0] & EDGE ) ? a.dstE == nullptr : a.dstV == nullptr )
      return;
   double          sum = 0.0;
   for ( long long e = e0; e < e1; e++ ) // ascending macro-cell id: every run adds in the same order
   {
      const long long off = a.grpEntry[e];
      sum += ( off & EDGE ) ? a.accE[off & ~EDGE] : a.accV[off];
   }
   const long long first = a.grpEntry[e0];
   const double    old   = ( first & EDGE ) ? a.dstE[first & ~EDGE] : a.dstV[first];
   const double    val   = ( a.beta == 0.0 ? 0.0 : a.beta * old ) + sum;
   for ( long long e = e0; e < e1; e++ )
   {
      const long long off = a.grpEntry[e];
      if ( off & EDGE )
         a.dstE[off & ~EDGE] = val;
      else
         a.dstV[off] = val;
   }
}

void hb_p2_launch_finalize( const HbP2View& pv, const HbP2FinalizeArgs& a, cudaStream_t st )
{
   const long long n = (long long) pv.nCells * ( (long long) pv.vCellSize + pv.eCellSize );
   if ( n > 0 )
      HB_LAUNCH( hb_p2_finalize_interior_kernel, (unsigned) ( ( n + 255 ) / 256 ), 256, st, pv, a );
   if ( a.nGroups > 0 )
      HB_LAUNCH( hb_p2_finalize_groups_kernel, (unsigned) ( ( a.nGroups + 127 ) / 128 ), 128, st, pv, a );
}

// ---- BLAS-1 on the edge block -------------------------------------------------------------------------------------------------------

__global__ void hb_p2_edge_lincomb_kernel( const HbP2View pv, const HbP2LinComb lc )
{
   const long long n = (long long) pv.nCells * pv.eCellSize;
   const long long i = (long long) blockIdx.x * blockDim.x + threadIdx.x;
   if ( i >= n )
      return;
   const int cell  = (int) ( i / pv.eCellSize );
   const int local = (int) ( i - (long long) cell * pv.eCellSize );
   if ( !( ( lc.selMask >> pv.cells[cell].region[pv.emask[local]].bflag ) & 1 ) )
      return;
   double v;
   if ( lc.op == 1 )
   {
      v = 1.0;
      for ( int k = 0; k < lc.n; k++ )
         v *= lc.src[k][i];
   }
   else if ( lc.op == 2 )
      v = 1.0 / lc.dst[i];
   else
   {
      v = lc.shift + ( lc.accumulate ? lc.dst[i] : 0.0 );
      for ( int k = 0; k < lc.n; k++ )
         v += lc.c[k] * lc.src[k][i];
   }
   lc.dst[i] = v;
}

void hb_p2_launch_edge_lincomb( const HbP2View& pv, const HbP2LinComb& lc, cudaStream_t st )
{
   const long long n = (long long) pv.nCells * pv.eCellSize;
   if ( n > 0 )
      HB_LAUNCH( hb_p2_edge_lincomb_kernel, (unsigned) ( ( n + 255 ) / 256 ), 256, st, pv, lc );
}

// one partial sum per block, added sequentially inside the block by thread 0 of ... (emulation-friendly: every thread adds its own
// strided share into a per-thread slot, the host adds the slots): partial[ block * blockDim + thread ]
__global__ void hb_p2_edge_dot_kernel( const HbP2View pv, uint32_t selMask, const double* a, const double* b, double* partial )
{
   const long long n      = (long long) pv.nCells * pv.eCellSize;
   const long long stride = (long long) gridDim.x * blockDim.x;
   const long long t      = (long long) blockIdx.x * blockDim.x + threadIdx.x;
   double          s      = 0.0;
   for ( long long i = t; i < n; i += stride )
   {
      const int       cell  = (int) ( i / pv.eCellSize );
      const int       local = (int) ( i - (long long) cell * pv.eCellSize );
      const HbRegion  reg   = pv.cells[cell].region[pv.emask[local]];
      if ( ( reg.owner & 1 ) && ( ( selMask >> reg.bflag ) & 1 ) )
         s += a[i] * b[i];
   }
   partial[t] = s;
}

int hb_p2_launch_edge_dot( const HbP2View& pv, uint32_t selMask, const double* a, const double* b, double* partial, int maxBlocks, cudaStream_t st )
{
   const long long n      = (long long) pv.nCells * pv.eCellSize;
   int             blocks = (int) std::min< long long >( maxBlocks, ( n + 4095 ) / 4096 );
   if ( blocks < 1 )
      blocks = 1;
   HB_LAUNCH( hb_p2_edge_dot_kernel, blocks, 64, st, pv, selMask, a, b, partial );
   return blocks * 64;
}

// ---- helpers for host callbacks and tests ---------------------------------------------------------------------------------------------

__global__ void hb_p2_quad_points_kernel( const HbP2View pv, const int form, double* out )
{
   const int perCell = ( pv.nAnchors + blockDim.x - 1 ) / blockDim.x;
   const int cell    = blockIdx.x / perCell;
   const int ai      = ( blockIdx.x - cell * perCell ) * blockDim.x + threadIdx.x;
   if ( cell >= pv.nCells || ai >= pv.nAnchors )
      return;
   const int          N  = pv.N;
   const HbBndPoint   p  = pv.anchors[ai];
   const HbP2CellGeom g  = pv.geom[cell];
   const int          nq = hb_p2_form_q2( form ) ? 4 : 11;
   for ( int t = 0; t < 6; t++ )
   {
      double* o = out + ( ( (size_t) cell * 6 + t ) * pv.nAnchors + ai ) * nq * 3;
      if ( p.x + p.y + p.z > N - hb_p2_slack[t] )
      {
         for ( int k = 0; k < nq * 3; k++ )
            o[k] = 0.0;
         continue;
      }
      double X[4][3];
      for ( int k = 0; k < 4; k++ )
         hb_p2_vertex_pos( g, 1.0 / N, p.x + hb_p2_vert[t][k][0], p.y + hb_p2_vert[t][k][1], p.z + hb_p2_vert[t][k][2], X[k] );
      for ( int q = 0; q < nq; q++ )
      {
         const double* Q = hb_p2_form_q2( form ) ? hb_p2_q2[q] : hb_p2_q4[q];
         double        xq[3], F[3], DF[9];
         for ( int d = 0; d < 3; d++ )
            xq[d] = Q[0] * X[0][d] + Q[1] * X[1][d] + Q[2] * X[2][d] + Q[3] * X[3][d];
         if ( pv.blending )
            hb_shell_map( g, xq, F, DF );
         for ( int d = 0; d < 3; d++ )
            o[3 * q + d] = pv.blending ? F[d] : xq[d];
      }
   }
}

void hb_p2_launch_quad_points( const HbP2View& pv, int form, double* out, cudaStream_t st )
{
   const int block   = 128;
   const int perCell = ( pv.nAnchors + block - 1 ) / block;
   if ( pv.nCells * perCell > 0 )
      HB_LAUNCH( hb_p2_quad_points_kernel, pv.nCells * perCell, block, st, pv, form, out );
}

__global__ void hb_p2_map_eval_kernel( const HbP2View pv, const int cell, const double* x, const int n, double* out )
{
   const int i = blockIdx.x * blockDim.x + threadIdx.x;
   if ( i >= n )
      return;
   const HbP2CellGeom g = pv.geom[cell];
   double             F[3], DF[9];
   const double       xx[3] = { x[3 * i], x[3 * i + 1], x[3 * i + 2] };
   hb_shell_map( g, xx, F, DF );
   for ( int d = 0; d < 3; d++ )
      out[12 * i + d] = F[d];
   for ( int d = 0; d < 9; d++ )
      out[12 * i + 3 + d] = DF[d];
}

void hb_p2_launch_map_eval( const HbP2View& pv, int cell, const double* x, int n, double* out, cudaStream_t st )
{
   if ( n > 0 )
      HB_LAUNCH( hb_p2_map_eval_kernel, ( n + 63 ) / 64, 64, st, pv, cell, x, n, out );
}

// ---- quadratic grid transfer -------------------------------------------------------------------------------------------------------------
// P2toP2QuadraticProlongation is nodal interpolation: a fine DoF takes the value of the coarse P2 function at its position; the
// restriction is the transpose (reference src/hyteg/gridtransferoperators/P2toP2QuadraticProlongation.cpp,
// P2toP2QuadraticRestriction.cpp:131-251 and their generated macro-cell kernels; pinned in tests/golden/ref_p2.npz).  Instead of the
// reference's per-orientation stencil kernels every fine DoF locates ONE coarse micro-tet that contains it by integer arithmetic:
// in units of a quarter of the coarse mesh width a fine vertex (x,y,z) sits at q = 2 (x,y,z) and a fine edge DoF at the sum of its
// two end points; inside the coarse cube at anchor a the six micro-cell types are unimodular, so 4 lambda (barycentric coordinates
// times four) are integers obtained with the integer inverse below.

// inverse of [ V1 - V0, V2 - V0, V3 - V0 ] (columns) per micro-cell type
HB_CONSTANT int8_t hb_p2_inv[6][3][3] = { { { 1, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 } },  { { 1, 1, 0 }, { -1, 0, 0 }, { 0, 0, 1 } }, { { 0, 1, 0 }, { 1, 1, 1 }, { -1, -1, 0 } },
                                          { { 1, 1, 1 }, { -1, 0, 0 }, { 0, -1, 0 } }, { { 0, 1, 1 }, { -1, -1, 0 }, { 0, 0, -1 } }, { { 1, 1, 0 }, { 0, -1, 0 }, { 0, 1, 1 } } };

// position of a fine DoF (quarter units of the coarse width) -> a coarse micro-tet containing it: anchor (ax,ay,az), type t, l4 = 4 lambda
__host__ __device__ __forceinline__ bool hb_p2_locate( int Nc, const int q[3], int& ax, int& ay, int& az, int& t, int l4[4] )
{
   for ( int cand = 0; cand < 8; cand++ )
   {
      ax = ( q[0] >> 2 ) - ( cand & 1 );
      ay = ( q[1] >> 2 ) - ( ( cand >> 1 ) & 1 );
      az = ( q[2] >> 2 ) - ( cand >> 2 );
      if ( ax < 0 || ay < 0 || az < 0 )
         continue;
      const int r[3] = { q[0] - 4 * ax, q[1] - 4 * ay, q[2] - 4 * az };
      if ( r[0] > 4 || r[1] > 4 || r[2] > 4 )
         continue;
      for ( t = 0; t < 6; t++ )
      {
         if ( ax + ay + az > Nc - hb_p2_slack[t] )
            continue;
         const int d[3] = { r[0] - 4 * hb_p2_vert[t][0][0], r[1] - 4 * hb_p2_vert[t][0][1], r[2] - 4 * hb_p2_vert[t][0][2] };
         bool      ok   = true;
         int       sum  = 0;
         for ( int k = 0; k < 3; k++ )
         {
            l4[k + 1] = hb_p2_inv[t][k][0] * d[0] + hb_p2_inv[t][k][1] * d[1] + hb_p2_inv[t][k][2] * d[2];
            ok        = ok && l4[k + 1] >= 0;
            sum += l4[k + 1];
         }
         l4[0] = 4 - sum;
         if ( ok && l4[0] >= 0 )
            return true;
      }
   }
   return false;
}

// the ten P2 basis functions at lambda = l4 / 4
__host__ __device__ __forceinline__ void hb_p2_basis( const int l4[4], double phi[10] )
{
   double lam[4];
   for ( int k = 0; k < 4; k++ )
   {
      lam[k] = 0.25 * l4[k];
      phi[k] = lam[k] * ( 2.0 * lam[k] - 1.0 );
   }
   for ( int k = 0; k < 6; k++ )
      phi[4 + k] = 4.0 * lam[hb_p2_pair[k][0]] * lam[hb_p2_pair[k][1]];
}

// quarter-unit position of entry 'local' of the fine vertex array (edge = false) or edge block (edge = true); false for padding
__host__ __device__ __forceinline__ bool hb_p2_fine_position( const HbP2View& fine, bool edge, int local, const HbBndPoint* vxyz, int q[3] )
{
   if ( !edge )
   {
      const HbBndPoint p = vxyz[local];
      q[0]               = 2 * p.x;
      q[1]               = 2 * p.y;
      q[2]               = 2 * p.z;
      return true;
   }
   // edge block: orientation o, then index inside the tetrahedral array of width w: recover (x,y,z) by search over z, y
   const int N = fine.N, T = hb_p2_tet( N );
   const int o = local < 6 * T ? local / T : 6;
   int       i = local - o * T;
   const int w = o < 6 ? N : N - 1;
   int       z = 0;
   while ( hb_p2_tet( w ) - hb_p2_tet( w - z - 1 ) <= i )
      z++;
   i -= hb_p2_tet( w ) - hb_p2_tet( w - z );
   const int wz = w - z;
   int       y  = 0;
   while ( ( y + 1 ) * ( wz + 1 ) - ( ( y + 1 ) * ( y + 2 ) ) / 2 <= i )
      y++;
   const int x = i - ( y * ( wz + 1 ) - ( y * ( y + 1 ) ) / 2 );
   q[0]        = 2 * x + hb_p2_ends[o][0][0] + hb_p2_ends[o][1][0];
   q[1]        = 2 * y + hb_p2_ends[o][0][1] + hb_p2_ends[o][1][1];
   q[2]        = 2 * z + hb_p2_ends[o][0][2] + hb_p2_ends[o][1][2];
   return true;
}

struct HbP2CoarseDofs
{
   int v[4], e[6];
};
__host__ __device__ __forceinline__ void hb_p2_coarse_dofs( int Nc, int ax, int ay, int az, int t, HbP2CoarseDofs& d )
{
   for ( int k = 0; k < 4; k++ )
      d.v[k] = hb_index( Nc, ax + hb_p2_vert[t][k][0], ay + hb_p2_vert[t][k][1], az + hb_p2_vert[t][k][2] );
   for ( int k = 0; k < 6; k++ )
      d.e[k] = hb_edge_index( Nc, hb_p2_edge[t][k][0], ax + hb_p2_edge[t][k][1], ay + hb_p2_edge[t][k][2], az + hb_p2_edge[t][k][3] );
}

template < bool RESTRICT >
__global__ void hb_p2_transfer_kernel( const HbP2View coarse, const HbP2View fine, const HbP2TransferArgs a )
{
   const long long nV = (long long) fine.nCells * fine.vCellSize, nE = (long long) fine.nCells * fine.eCellSize;
   const long long i  = (long long) blockIdx.x * blockDim.x + threadIdx.x;
   if ( i >= nV + nE )
      return;
   const bool      edge  = i >= nV;
   const long long j     = edge ? i - nV : i;
   const int       csz   = edge ? fine.eCellSize : fine.vCellSize;
   const int       cell  = (int) ( j / csz );
   const int       local = (int) ( j - (long long) cell * csz );
   const int       mask  = edge ? fine.emask[local] : fine.vmask[local];
   const HbRegion  reg   = fine.cells[cell].region[mask];
   if ( RESTRICT )
   {
      if ( !( reg.owner & 1 ) )
         return; // every fine DoF contributes once
   }
   else if ( !( ( a.selMask >> reg.bflag ) & 1 ) )
      return;
   int q[3], ax, ay, az, t, l4[4];
   hb_p2_fine_position( fine, edge, local, a.fineXyz, q );
   if ( !hb_p2_locate( coarse.N, q, ax, ay, az, t, l4 ) )
      return;
   double phi[10];
   hb_p2_basis( l4, phi );
   HbP2CoarseDofs d;
   hb_p2_coarse_dofs( coarse.N, ax, ay, az, t, d );
   const size_t vb = (size_t) cell * coarse.vCellSize, eb = (size_t) cell * coarse.eCellSize;
   if ( RESTRICT )
   {
      const double val = edge ? a.fineE[j] : a.fineV[j];
      for ( int k = 0; k < 4; k++ )
         if ( phi[k] != 0.0 )
            hb_atomic_add( a.coarseV + vb + d.v[k], phi[k] * val );
      for ( int k = 0; k < 6; k++ )
         if ( phi[4 + k] != 0.0 )
            hb_atomic_add( a.coarseE + eb + d.e[k], phi[4 + k] * val );
   }
   else
   {
      double val = 0.0;
      for ( int k = 0; k < 4; k++ )
         if ( phi[k] != 0.0 )
            val += phi[k] * a.coarseV[vb + d.v[k]];
      for ( int k = 0; k < 6; k++ )
         if ( phi[4 + k] != 0.0 )
            val += phi[4 + k] * a.coarseE[eb + d.e[k]];
      double* dst = const_cast< double* >( edge ? a.fineE : a.fineV );
      dst[j]      = a.add ? dst[j] + val : val;
   }
}

void hb_p2_launch_prolongate( const HbP2View& coarse, const HbP2View& fine, const HbP2TransferArgs& a, cudaStream_t st )
{
   const long long n = (long long) fine.nCells * ( (long long) fine.vCellSize + fine.eCellSize );
   if ( n > 0 )
      HB_LAUNCH( hb_p2_transfer_kernel< false >, (unsigned) ( ( n + 127 ) / 128 ), 128, st, coarse, fine, a );
}

void hb_p2_launch_restrict( const HbP2View& coarse, const HbP2View& fine, const HbP2TransferArgs& a, cudaStream_t st )
{
   const long long n = (long long) fine.nCells * ( (long long) fine.vCellSize + fine.eCellSize );
   if ( n > 0 )
      HB_LAUNCH( hb_p2_transfer_kernel< true >, (unsigned) ( ( n + 127 ) / 128 ), 128, st, coarse, fine, a );
}
