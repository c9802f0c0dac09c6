#include <stdio.h>
#include <stdlib.h>
void mrref_stub_0(void) __asm__("_ZN2MR11MeshBuilder16VertexIdentifier12addTrianglesERKSt6vectorISt5arrayINS_7Vector3IfEELm3EESaIS6_EE");
void mrref_stub_0(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR11MeshBuilder16VertexIdentifier12addTrianglesERKSt6vectorISt5arrayINS_7Vector3IfEELm3EESaIS6_EE\n"); abort(); }
void mrref_stub_1(void) __asm__("_ZN2MR11MeshBuilder16VertexIdentifier7reserveEm");
void mrref_stub_1(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR11MeshBuilder16VertexIdentifier7reserveEm\n"); abort(); }
void mrref_stub_2(void) __asm__("_ZN2MR11MeshBuilder28duplicateNonManifoldVerticesERNS_6VectorISt5arrayINS_2IdINS_7VertTagEEELm3EENS3_INS_7FaceTagEEEEEPNS_11TypedBitSetIS8_EEPSt6vectorINS0_15VertDuplicationESaISF_EES5_RKSt8functionIFbS5_S5_S5_S5_S5_EE");
void mrref_stub_2(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR11MeshBuilder28duplicateNonManifoldVerticesERNS_6VectorISt5arrayINS_2IdINS_7VertTagEEELm3EENS3_INS_7FaceTagEEEEEPNS_11TypedBitSetIS8_EEPSt6vectorINS0_15VertDuplicationESaISF_EES5_RKSt8functionIFbS5_S5_S5_S5_S5_EE\n"); abort(); }
void mrref_stub_3(void) __asm__("_ZN2MR11Src2TgtMapsC1EPNS_6VectorINS_2IdINS_7FaceTagEEES4_EEPNS1_INS2_INS_7VertTagEEES8_EEPNS1_INS2_INS_7EdgeTagEEENS2_INS_17UndirectedEdgeTagEEEEE");
void mrref_stub_3(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR11Src2TgtMapsC1EPNS_6VectorINS_2IdINS_7FaceTagEEES4_EEPNS1_INS2_INS_7VertTagEEES8_EEPNS1_INS2_INS_7EdgeTagEEENS2_INS_17UndirectedEdgeTagEEEEE\n"); abort(); }
void mrref_stub_4(void) __asm__("_ZN2MR11Src2TgtMapsD1Ev");
void mrref_stub_4(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR11Src2TgtMapsD1Ev\n"); abort(); }
void mrref_stub_5(void) __asm__("_ZN2MR12readByBlocksERSiPcmSt8functionIFbfEEm");
void mrref_stub_5(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR12readByBlocksERSiPcmSt8functionIFbfEEm\n"); abort(); }
void mrref_stub_6(void) __asm__("_ZN2MR13getStreamSizeERSi");
void mrref_stub_6(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR13getStreamSizeERSi\n"); abort(); }
void mrref_stub_7(void) __asm__("_ZN2MR13subdivideMeshERNS_4MeshERKNS_17SubdivideSettingsE");
void mrref_stub_7(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR13subdivideMeshERNS_4MeshERKNS_17SubdivideSettingsE\n"); abort(); }
void mrref_stub_8(void) __asm__("_ZN2MR14AABBTreePoints5refitERKNS_6VectorINS_7Vector3IfEENS_2IdINS_7VertTagEEEEERKNS_11TypedBitSetIS6_EE");
void mrref_stub_8(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR14AABBTreePoints5refitERKNS_6VectorINS_7Vector3IfEENS_2IdINS_7VertTagEEEEERKNS_11TypedBitSetIS6_EE\n"); abort(); }
void mrref_stub_9(void) __asm__("_ZN2MR14AABBTreePointsC1ERKNS_4MeshE");
void mrref_stub_9(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR14AABBTreePointsC1ERKNS_4MeshE\n"); abort(); }
void mrref_stub_10(void) __asm__("_ZN2MR15getEdgeOrderingERKNS_4BMapINS_2IdINS_7FaceTagEEES3_EERKNS_12MeshTopologyE");
void mrref_stub_10(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR15getEdgeOrderingERKNS_4BMapINS_2IdINS_7FaceTagEEES3_EERKNS_12MeshTopologyE\n"); abort(); }
void mrref_stub_11(void) __asm__("_ZN2MR17getVertexOrderingERKNS_4BMapINS_2IdINS_7FaceTagEEES3_EERKNS_12MeshTopologyE");
void mrref_stub_11(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR17getVertexOrderingERKNS_4BMapINS_2IdINS_7FaceTagEEES3_EERKNS_12MeshTopologyE\n"); abort(); }
void mrref_stub_12(void) __asm__("_ZN2MR18computeBoundingBoxINS_7Vector3IfEEEENS_3BoxIT_EERKNS_6VectorIS4_NS_2IdINS_7VertTagEEEEES9_S9_PKNS_11TypedBitSetIS9_EEPKNS_8AffineXfIS4_EE");
void mrref_stub_12(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR18computeBoundingBoxINS_7Vector3IfEEEENS_3BoxIT_EERKNS_6VectorIS4_NS_2IdINS_7VertTagEEEEES9_S9_PKNS_11TypedBitSetIS9_EEPKNS_8AffineXfIS4_EE\n"); abort(); }
void mrref_stub_13(void) __asm__("_ZN2MR18splitOnSimpleLoopsERKNS_12MeshTopologyEOSt6vectorIS3_INS_2IdINS_7EdgeTagEEESaIS6_EESaIS8_EE");
void mrref_stub_13(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR18splitOnSimpleLoopsERKNS_12MeshTopologyEOSt6vectorIS3_INS_2IdINS_7EdgeTagEEESaIS6_EESaIS8_EE\n"); abort(); }
void mrref_stub_14(void) __asm__("_ZN2MR19executeHoleFillPlanERNS_4MeshENS_2IdINS_7EdgeTagEEERNS_12HoleFillPlanEPNS_11TypedBitSetINS2_INS_7FaceTagEEEEE");
void mrref_stub_14(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR19executeHoleFillPlanERNS_4MeshENS_2IdINS_7EdgeTagEEERNS_12HoleFillPlanEPNS_11TypedBitSetINS2_INS_7FaceTagEEEEE\n"); abort(); }
void mrref_stub_15(void) __asm__("_ZN2MR22getOptimalFaceOrderingERKNS_4MeshE");
void mrref_stub_15(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR22getOptimalFaceOrderingERKNS_4MeshE\n"); abort(); }
void mrref_stub_16(void) __asm__("_ZN2MR22getPlanarHoleFillPlansERKNS_4MeshERKSt6vectorINS_2IdINS_7EdgeTagEEESaIS6_EE");
void mrref_stub_16(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR22getPlanarHoleFillPlansERKNS_4MeshERKSt6vectorINS_2IdINS_7EdgeTagEEESaIS6_EE\n"); abort(); }
void mrref_stub_17(void) __asm__("_ZN2MR25findSmallestCloseVerticesERKNS_4MeshEfRKSt8functionIFbfEE");
void mrref_stub_17(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR25findSmallestCloseVerticesERKNS_4MeshEfRKSt8functionIFbfEE\n"); abort(); }
void mrref_stub_18(void) __asm__("_ZN2MR25findSmallestCloseVerticesERKNS_6VectorINS_7Vector3IfEENS_2IdINS_7VertTagEEEEEfPKNS_11TypedBitSetIS5_EERKSt8functionIFbfEE");
void mrref_stub_18(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR25findSmallestCloseVerticesERKNS_6VectorINS_7Vector3IfEENS_2IdINS_7VertTagEEEEEfPKNS_11TypedBitSetIS5_EERKSt8functionIFbfEE\n"); abort(); }
void mrref_stub_19(void) __asm__("_ZN2MR4sameERKNS_12MeshTopologyERKNS_9EdgePointES5_");
void mrref_stub_19(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR4sameERKNS_12MeshTopologyERKNS_9EdgePointES5_\n"); abort(); }
void mrref_stub_20(void) __asm__("_ZN2MR7reverseERSt6vectorINS_2IdINS_7EdgeTagEEESaIS3_EE");
void mrref_stub_20(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR7reverseERSt6vectorINS_2IdINS_7EdgeTagEEESaIS3_EE\n"); abort(); }
void mrref_stub_21(void) __asm__("_ZN2MR8makeCubeERKNS_7Vector3IfEES3_");
void mrref_stub_21(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR8makeCubeERKNS_7Vector3IfEES3_\n"); abort(); }
void mrref_stub_22(void) __asm__("_ZN2MR9EdgePointC1ERKNS_12MeshTopologyENS_2IdINS_7VertTagEEE");
void mrref_stub_22(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZN2MR9EdgePointC1ERKNS_12MeshTopologyENS_2IdINS_7VertTagEEE\n"); abort(); }
void mrref_stub_23(void) __asm__("_ZNK2MR14AABBTreePoints9heapBytesEv");
void mrref_stub_23(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZNK2MR14AABBTreePoints9heapBytesEv\n"); abort(); }
void mrref_stub_24(void) __asm__("_ZNK2MR9FloatGridcvbEv");
void mrref_stub_24(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZNK2MR9FloatGridcvbEv\n"); abort(); }
void mrref_stub_25(void) __asm__("_ZNK2MR9FloatGridptEv");
void mrref_stub_25(void) { fprintf(stderr, "oracle/_ref: unlinked reference symbol called: _ZNK2MR9FloatGridptEv\n"); abort(); }
