ssh test@127.0.0.1 "kill -9 9319"
ssh test@127.0.0.1 "kill -9 9320"
ssh test@127.0.0.1 "kill -9 9321"
ssh test@127.0.0.1 "kill -9 9322"
kill -9 8873
