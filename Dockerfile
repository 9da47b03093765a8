# garfield-b200: build image. Counterpart of the reference's pytorch_impl/Dockerfile (which pip-installs the requirements and
# runs `quart init_demo` so that the native modules are compiled into the image). Here the image compiles both native
# libraries in tree — the host library with g++, every CUDA source for sm_100a with nvcc (cross-compiles without a GPU) —
# runs the CPU test-suite, and serves the LEARN demo by default.
#
#   docker build -t garfield-b200 .
#   docker run --rm -p 8080:8080 garfield-b200                                   # LEARN web demo on :8080
#   docker run --rm --gpus all garfield-b200 python bench.py --gpus 1            # headline benchmark on a B200
ARG BASE=nvcr.io/nvidia/pytorch:25.04-py3
FROM ${BASE}
WORKDIR /opt/garfield-b200
COPY . .
ENV GARFIELD_QUIET=1 PYTHONUNBUFFERED=1
# build() = g++ host library + nvcc -gencode arch=compute_100a,code=sm_100a for every .cu + import of the package
RUN python -c "import __graft_entry__ as g; print(g.build.__doc__); from garfield_b200 import _build; print(_build.build_all())"
RUN python -m pytest tests -x -q -m "not gpu" -p no:cacheprovider
RUN python applications/LEARN/demo.py --init_demo
EXPOSE 8080
CMD ["python", "applications/LEARN/demo.py", "--host", "0.0.0.0", "--port", "8080", "--transport", "processes"]
