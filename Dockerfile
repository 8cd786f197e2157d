# Container image for the hola service.
#
# Counterpart of /root/reference/Dockerfile (base image + ENV + EXPOSE 8080 + ADD of the uber-jar)
# and of the inline image definition in /root/reference/pom.xml:247-280. The Java image needs a
# JDK base layer; this one is a single static executable on an empty file system.
#
# NOT BUILT OR RUN on the development box: there is no container runtime there and no network to
# pull a toolchain image (BASELINE.md section 2). The static executable itself is built and
# tested outside a container (`make static`, `pytest --variant static`).
#
#   docker build -t redhatmsa/hola:0.0.1-SNAPSHOT .
#   docker run -p 8080:8080 redhatmsa/hola:0.0.1-SNAPSHOT

FROM gcc:13 AS build
WORKDIR /src
COPY Makefile ./
COPY include ./include
COPY src ./src
COPY tests/cpp ./tests/cpp
RUN make -j"$(nproc)" static EXTRA_LDFLAGS=-s && ./build/hola-static --version

FROM scratch
# Same variable, same default as the reference image (Dockerfile:6, pom.xml:271). Unlike the
# reference, the service also starts when it is unset (tracing is then off).
ENV ZIPKIN_SERVER_URL=http://zipkin-query:9411
# The reference image also sets JAVA_APP_JAR, AB_OFF (Jolokia agent off) and JAVA_OPTIONS=-Xmx512m
# (Dockerfile:3-5). They configure a JVM and have no meaning here; the manifest in deploy/ keeps
# AB_OFF and JAVA_OPTIONS only so that tooling which expects them keeps working.
EXPOSE 8080
COPY --from=build /src/build/hola-static /app/hola
# HOSTNAME is set by the container runtime to the container id / pod name, which is what
# GET /api/hola reports (HolaResource.java:57).
ENTRYPOINT ["/app/hola"]
