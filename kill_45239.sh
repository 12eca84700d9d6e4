ssh test@127.0.0.1 "kill -9 3353"
ssh test@127.0.0.1 "kill -9 3354"
ssh test@127.0.0.1 "kill -9 3355"
kill -9 2946
