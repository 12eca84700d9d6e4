ssh test@127.0.0.1 "kill -9 3249"
ssh test@127.0.0.1 "kill -9 3250"
kill -9 2946
