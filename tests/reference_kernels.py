"""Kernel strings of the reference, verbatim OpenCL C — TEST VECTORS ONLY.

They come from the reference's own tests and sample (Tester.cs:35-38, :117-121, :7727-7742, :7820-7824; Kamera.cs:234-245)
and are kept here, outside the product package, because what the tests prove with them is that unmodified reference
kernels build and run through libKutuphaneCL.so.  Nothing under cekirdekler_b200/ uses them.
"""

def reference_copy_kernel(ctype: str) -> str:
    """The `test` copy kernel of the reference's Tester matrix for one element type (Tester.cs:35-38, :87-91 ...)."""
    return (
        "__kernel void test(__global " + ctype + " * data,__global " + ctype + " * data2)\n"
        "{\n"
        "    int id=get_global_id(0);\n"
        "    data2[id]=data[id];\n"
        "}\n"
    )


# C# char (2 bytes) is copied as two uchars per work-item (Tester.cs:117-121)
REFERENCE_COPY_KERNEL_CHAR = """
                __kernel void test(__global uchar * data,__global uchar * data2)
                { 
                    data2[get_global_id(0)*2]=data[get_global_id(0)*2];  
                    data2[get_global_id(0)*2+1]=data[get_global_id(0)*2+1];  
                }"""

REFERENCE_VECTOR_ADD = """
                __kernel void vectorAdd(__global float*a,__global float*b,__global float*c)
                {
                    int i=get_global_id(0);
                    c[i]=a[i]+b[i];
                }"""


def reference_nbody_kernel(n: int) -> str:
    """The reference's nBody kernel text with n spliced in, as Tester.cs:7727-7742 builds it."""
    return """
                __kernel void nBody(__global float * x,__global float * y,__global float * fx,__global float * fy)
                {
                    int i=get_global_id(0);    
                    float fx0 = 0;
                    float fy0 = 0;
                    for (int j = 0; j < """ + str(n) + """; j++)
                    {
                        float dx = x[i] - x[j];
                        float dy = y[i] - y[j];
                        float r = (float)sqrt(dx*dx+dy*dy+0.0001f);
                        fx0 += (dx * (1.0f / (r * r * r)));
                        fy0 += (dx * (1.0f / (r * r * r)));
                    }
                    fx[i] = fx0;
                    fy[i] = fy0;
                }
                """


# The Unity sample's kernel, verbatim (/root/reference/Kamera.cs:234-245): Vector3[] arrays arrive as byte arrays wrapped with
# ClArray<byte>.wrapArrayOfStructs (numberOfElementsPerWorkItem = sizeof(Vector3) = 12) and are read as floats by the kernel.
REFERENCE_WAVE_EQUATION = """
                        __kernel void waveEquation( __global float *xyz,__global float *xyzn,__global float *xyzo,__global float * arguments)
                        {
                            int threadId=get_global_id(0);
                            if(threadId<arguments[4])
                            {
                                float dx=xyz[threadId*3]-arguments[2];float dy=xyz[threadId*3+1]-arguments[3];float t=arguments[1];
                                float ctr=arguments[0];float wave=0.02f*ctr*sin(40.0f*t+100.0f*sqrt(dx*dx+dy*dy));
                                xyzo[threadId*3]=xyz[threadId*3]+xyzn[threadId*3]*wave; // wave equation for all surface vertices
                                xyzo[threadId*3+1]=xyz[threadId*3+1]+xyzn[threadId*3+1]*wave; // wave equation for all surface vertices
                                xyzo[threadId*3+2]=xyz[threadId*3+2]+xyzn[threadId*3+2]*wave; // wave equation for all surface vertices
                            }
                        }
                    """
