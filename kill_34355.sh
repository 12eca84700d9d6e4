ssh test@127.0.0.1 "kill -9 3035"
ssh test@127.0.0.1 "kill -9 3036"
ssh test@127.0.0.1 "kill -9 3037"
ssh test@127.0.0.1 "kill -9 3038"
kill -9 2946
